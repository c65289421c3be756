// Headless stand-in for the reference's detect_grasps_file executable (src/nodes/detect_grasps_file.cpp:6-32 +
// GraspDetectionNode::detectGraspPosesInFile, src/nodes/grasp_detection_node.cpp:98-120) on top of the header-only
// shims: load one .pcd (or <prefix>l_reg.pcd + <prefix>r_reg.pcd), preprocessPointCloud, detectGraspPoses, print the
// grasps.  The ROS parameter server is replaced by `name:=value` arguments with the parameter names of
// launch/file_detect_grasps.launch; what that launch file sets is the default here.
//
//   detect_grasps_file cloud_file_name:=scene.pcd [num_samples:=5000] [antipodal_mode:=1] [trained_file:=x.caffemodel]
//                      [use_importance_sampling:=true] [num_gpus:=8] [hands_file:=grasps.csv] [msg_file:=grasps.bin] ...
//
// Build: make -C examples   (needs libag2.so: python -c "import __graft_entry__ as g; g.build()").
#include <agile_grasp2/cloud_camera.h>
#include <agile_grasp2/grasp_detector.h>
#include <agile_grasp2/importance_sampling.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

namespace {

typedef std::map<std::string, std::string> Args;

bool get(const Args& a, const char* k, std::string& v) {
  Args::const_iterator it = a.find(k);
  if (it == a.end()) return false;
  v = it->second;
  return true;
}
void param(const Args& a, const char* k, int& v) {
  std::string s;
  if (get(a, k, s)) v = std::atoi(s.c_str());
}
void param(const Args& a, const char* k, double& v) {
  std::string s;
  if (get(a, k, s)) v = std::atof(s.c_str());
}
void param(const Args& a, const char* k, uint64_t& v) {
  std::string s;
  if (get(a, k, s)) v = std::strtoull(s.c_str(), nullptr, 10);
}
void param(const Args& a, const char* k, bool& v) {
  std::string s;
  if (get(a, k, s)) v = (s == "true" || s == "1" || s == "True");
}
void param(const Args& a, const char* k, std::string& v) { get(a, k, v); }
// "[a, b, c]" or "a,b,c"
void param(const Args& a, const char* k, std::vector<double>& v) {
  std::string s;
  if (!get(a, k, s)) return;
  v.clear();
  const char* p = s.c_str();
  while (*p) {
    char* end = nullptr;
    double x = std::strtod(p, &end);
    if (end == p) {
      p++;
      continue;
    }
    v.push_back(x);
    p = end;
  }
}

}  // namespace

int main(int argc, char** argv) {
  Args args;
  for (int i = 1; i < argc; i++) {
    std::string s = argv[i];
    size_t at = s.find(":=");
    if (at == std::string::npos) {
      std::fprintf(stderr, "argument '%s' is not name:=value\n", argv[i]);
      return 2;
    }
    size_t from = s[0] == '_' ? 1 : 0;  // rosrun's private-parameter form _name:=value
    args[s.substr(from, at - from)] = s.substr(at + 2);
  }
  std::string cloud_file_name;
  if (!get(args, "cloud_file_name", cloud_file_name)) {
    std::fprintf(stderr, "usage: %s cloud_file_name:=<file.pcd | prefix of l_reg.pcd / r_reg.pcd> [name:=value ...]\n", argv[0]);
    return 2;
  }

  // launch/file_detect_grasps.launch:14-48
  GraspDetector::Params prm;
  prm.workspace = {0.48, 0.92, -0.6, 0.4, -0.25, 1.0};
  prm.num_samples = 5000;
  prm.num_threads = 4;
  prm.antipodal_mode = AG2_MODE_PREDICTION;
  prm.filter_half_grasps = false;
  prm.gripper_width_range = {0.03, 0.08};
  prm.init_bite = 0.01;
  prm.min_score_diff = 300;
  prm.min_inliers = 5;
  prm.num_selected = 30;
  param(args, "workspace", prm.workspace);
  param(args, "camera_pose", prm.camera_pose);
  param(args, "num_samples", prm.num_samples);
  param(args, "num_threads", prm.num_threads);
  param(args, "nn_radius_taubin", prm.nn_radius_taubin);
  param(args, "nn_radius_hands", prm.nn_radius_hands);
  param(args, "num_orientations", prm.num_orientations);
  param(args, "antipodal_mode", prm.antipodal_mode);
  param(args, "voxelize", prm.voxelize);
  param(args, "filter_half_grasps", prm.filter_half_grasps);
  param(args, "gripper_width_range", prm.gripper_width_range);
  param(args, "finger_width", prm.finger_width);
  param(args, "hand_outer_diameter", prm.hand_outer_diameter);
  param(args, "hand_depth", prm.hand_depth);
  param(args, "hand_height", prm.hand_height);
  param(args, "init_bite", prm.init_bite);
  param(args, "model_file", prm.model_file);
  param(args, "trained_file", prm.trained_file);  // empty: the prototxt's Xavier fillers (scores are then synthetic)
  param(args, "label_file", prm.label_file);
  param(args, "min_score_diff", prm.min_score_diff);
  param(args, "min_inliers", prm.min_inliers);
  param(args, "num_selected", prm.num_selected);
  param(args, "rng_seed", prm.rng_seed);  // new: the reference seeds from the clock
  param(args, "num_gpus", prm.num_gpus);  // new: GPUs of this box to shard the samples over
  bool use_importance_sampling = false;
  param(args, "use_importance_sampling", use_importance_sampling);
  std::string hands_file, msg_file;
  param(args, "hands_file", hands_file);  // GraspHypothesis::writeHandsToFile
  param(args, "msg_file", msg_file);      // the bytes of the GraspListMsg the node would publish

  try {
    // src/nodes/detect_grasps_file.cpp:15-26
    std::string file_name_left, file_name_right;
    if (cloud_file_name.find(".pcd") == std::string::npos) {
      file_name_left = cloud_file_name + "l_reg.pcd";
      file_name_right = cloud_file_name + "r_reg.pcd";
    } else {
      file_name_left = cloud_file_name;
    }
    typedef std::chrono::steady_clock Clock;
    Clock::time_point t0 = Clock::now();
    // grasp_detection_node.cpp:98-120
    CloudCamera cloud_cam = file_name_right.empty() ? CloudCamera(file_name_left) : CloudCamera(file_name_left, file_name_right);
    std::printf("Loaded %zu points\n", cloud_cam.getCloudOriginal().size());
    if (cloud_cam.getCloudOriginal().size() == 0) {  // grasp_detector.cpp:86-91
      std::fprintf(stderr, "Point cloud is empty!\n");
      return 1;
    }
    Clock::time_point t1 = Clock::now();
    std::vector<GraspHypothesis> grasps;
    ag2_stats st;
    if (use_importance_sampling) {
      ImportanceSampling detector(prm);
      detector.preprocessPointCloud(cloud_cam);
      grasps = detector.detectGraspPoses(cloud_cam);
      st = detector.stats();
    } else {
      GraspDetector detector(prm);
      detector.preprocessPointCloud(cloud_cam);
      grasps = detector.detectGraspPoses(cloud_cam);
      st = detector.stats();
    }
    Clock::time_point t2 = Clock::now();
    std::printf("After preprocessing: %zu points, %zu samples\n", cloud_cam.getCloudProcessed().size(),
                cloud_cam.getSampleIndices().size());
    for (size_t i = 0; i < grasps.size(); i++) std::printf("grasp %zu, score: %g\n", i, grasps[i].getScore());  // grasp_detector.cpp:260-263
    std::printf("# grasps: %zu; load %.1f ms, create + preprocess + detect %.1f ms", grasps.size(),
                std::chrono::duration<double, std::milli>(t1 - t0).count(), std::chrono::duration<double, std::milli>(t2 - t1).count());
    if (prm.num_gpus <= 1)
      std::printf(" (device: normals %.2f, frames %.2f, hand search %.2f, images %.2f, classifier %.2f ms)", st.ms_phase[1],
                  st.ms_phase[2], st.ms_phase[3], st.ms_phase[5], st.ms_phase[6]);
    std::printf("\n");
    if (!hands_file.empty()) GraspHypothesis::writeHandsToFile(hands_file, grasps);
    if (!msg_file.empty()) {
      std::vector<uint8_t> wire = GraspHypothesis::toGraspListMsg(grasps, "base", 0, 0, 0);
      std::FILE* f = std::fopen(msg_file.c_str(), "wb");
      if (!f) throw std::runtime_error("cannot write " + msg_file);
      std::fwrite(wire.data(), 1, wire.size(), f);
      std::fclose(f);
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "detect_grasps_file: %s\n", e.what());
    return 1;
  }
  return 0;
}
