// Shim of the reference's GraspHypothesis (include/agile_grasp2/grasp_hypothesis.h:71-79,116-288):
// same getters / setters, backed by the 192-byte ag2_hypothesis record.
#ifndef AGILE_GRASP2_SHIM_GRASP_HYPOTHESIS_H_
#define AGILE_GRASP2_SHIM_GRASP_HYPOTHESIS_H_

#include <ag2.h>

#include <cstdio>
#include <string>
#include <vector>

#include "types.h"

class GraspHypothesis {
 public:
  GraspHypothesis() : origin_(nullptr) { rec_ = ag2_hypothesis(); }
  GraspHypothesis(const ag2_hypothesis& rec, const ag2_ctx* origin) : rec_(rec), origin_(origin) {}

  ag2shim::Vector3d getApproach() const { return ag2shim::Vector3d(rec_.approach); }
  ag2shim::Vector3d getAxis() const { return ag2shim::Vector3d(rec_.axis); }
  ag2shim::Vector3d getBinormal() const { return ag2shim::Vector3d(rec_.binormal); }
  ag2shim::Vector3d getGraspBottom() const { return ag2shim::Vector3d(rec_.bottom); }
  ag2shim::Vector3d getGraspSurface() const { return ag2shim::Vector3d(rec_.surface); }
  ag2shim::Vector3d getGraspTop() const { return ag2shim::Vector3d(rec_.top); }
  double getGraspWidth() const { return rec_.width; }
  double getScore() const { return rec_.score; }
  bool isHalfAntipodal() const { return rec_.half_antipodal != 0; }
  bool isFullAntipodal() const { return rec_.full_antipodal != 0; }

  void setGraspBottom(const ag2shim::Vector3d& v) { set(rec_.bottom, v); }
  void setGraspSurface(const ag2shim::Vector3d& v) { set(rec_.surface, v); }
  void setGraspTop(const ag2shim::Vector3d& v) { set(rec_.top, v); }
  void setGraspWidth(double w) { rec_.width = w; }
  void setScore(double s) { rec_.score = s; }
  void setHalfAntipodal(bool b) { rec_.half_antipodal = b ? 1 : 0; }
  void setFullAntipodal(bool b) { rec_.full_antipodal = b ? 1 : 0; }

  // GraspHypothesis::print, grasp_hypothesis.cpp:26-35
  void print() const {
    std::printf("axis: %g %g %g\napproach: %g %g %g\nbinormal: %g %g %g\ngrasp width: %g\n", rec_.axis[0],
                rec_.axis[1], rec_.axis[2], rec_.approach[0], rec_.approach[1], rec_.approach[2], rec_.binormal[0],
                rec_.binormal[1], rec_.binormal[2], rec_.width);
    std::printf("grasp surface: %g %g %g\ngrasp bottom: %g %g %g\ngrasp top: %g %g %g\n", rec_.surface[0],
                rec_.surface[1], rec_.surface[2], rec_.bottom[0], rec_.bottom[1], rec_.bottom[2], rec_.top[0],
                rec_.top[1], rec_.top[2]);
  }

  // agile_grasp2/GraspMsg (msg/GraspMsg.msg:5-16) as a plain struct: points and vectors in double, width and score
  // narrowed to float32 as std_msgs/Float32 does
  struct GraspMsg {
    struct Xyz {
      double x, y, z;
    };
    struct Float32 {
      float data;
    };
    Xyz surface, bottom, top, axis, approach, binormal;
    Float32 width, score;
  };
  // convertToGraspMsg, grasp_hypothesis.cpp:37-50
  GraspMsg convertToGraspMsg() const {
    GraspMsg m;
    m.surface = GraspMsg::Xyz{rec_.surface[0], rec_.surface[1], rec_.surface[2]};
    m.bottom = GraspMsg::Xyz{rec_.bottom[0], rec_.bottom[1], rec_.bottom[2]};
    m.top = GraspMsg::Xyz{rec_.top[0], rec_.top[1], rec_.top[2]};
    m.axis = GraspMsg::Xyz{rec_.axis[0], rec_.axis[1], rec_.axis[2]};
    m.approach = GraspMsg::Xyz{rec_.approach[0], rec_.approach[1], rec_.approach[2]};
    m.binormal = GraspMsg::Xyz{rec_.binormal[0], rec_.binormal[1], rec_.binormal[2]};
    m.width.data = (float)rec_.width;
    m.score.data = (float)rec_.score;
    return m;
  }

  // writeHandsToFile, grasp_hypothesis.cpp:4-23 (without its console echo): per hand one line of bottom, surface, axis,
  // approach, binormal -- each component followed by a comma, 17 significant digits as
  // boost::lexical_cast<std::string>(double) prints -- then the width as an ostream prints a double (%g)
  static void writeHandsToFile(const std::string& filename, const std::vector<GraspHypothesis>& hands) {
    std::FILE* f = std::fopen(filename.c_str(), "w");
    if (!f) return;  // the reference's ofstream fails silently too
    for (const GraspHypothesis& h : hands) {
      const double* v[5] = {h.rec_.bottom, h.rec_.surface, h.rec_.axis, h.rec_.approach, h.rec_.binormal};
      for (int k = 0; k < 5; k++) std::fprintf(f, "%.17g,%.17g,%.17g,", v[k][0], v[k][1], v[k][2]);
      std::fprintf(f, "%g\n", h.rec_.width);
    }
    std::fclose(f);
  }

  // convertToGraspMsg / createGraspListMsg (grasp_hypothesis.cpp:37-50, grasp_detection_node.cpp:306-313):
  // the ROS wire bytes of an agile_grasp2/GraspListMsg
  static std::vector<uint8_t> toGraspListMsg(const std::vector<GraspHypothesis>& hands, const std::string& frame_id,
                                             uint32_t seq, uint32_t sec, uint32_t nsec) {
    std::vector<ag2_hypothesis> rec(hands.size());
    for (size_t i = 0; i < hands.size(); i++) rec[i] = hands[i].rec_;
    size_t len = 0;
    ag2_serialize_grasp_list(rec.data(), rec.size(), seq, sec, nsec, frame_id.c_str(), nullptr, 0, &len);
    std::vector<uint8_t> out(len);
    ag2_serialize_grasp_list(rec.data(), rec.size(), seq, sec, nsec, frame_id.c_str(), out.data(), out.size(), &len);
    return out;
  }

  const ag2_hypothesis& record() const { return rec_; }
  const ag2_ctx* origin() const { return origin_; }

 private:
  static void set(double* dst, const ag2shim::Vector3d& v) {
    for (int a = 0; a < 3; a++) dst[a] = v(a);
  }
  ag2_hypothesis rec_;
  const ag2_ctx* origin_;  // the context whose last generate call owns the in-box points (for Learning)
};

#endif
