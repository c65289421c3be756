// Shims of the reference's Learning (include/agile_grasp2/learning.h:75-81, learning.cpp:4-33) and
// Classifier (include/agile_grasp2/caffe_classifier.h:55-66, caffe_classifier.cpp:4-127).
#ifndef AGILE_GRASP2_SHIM_LEARNING_H_
#define AGILE_GRASP2_SHIM_LEARNING_H_

#include <ag2.h>

#include <string>
#include <utility>
#include <vector>

#include "hand_search.h"

class Learning {
 public:
  Learning(int size, int num_threads) : size_(size), num_threads_(num_threads), own_(nullptr) {
    if (size != AG2_IMAGE_SIZE) throw std::runtime_error("Learning: only 60 x 60 grasp images are supported");
  }
  ~Learning() { ag2_destroy(own_); }
  Learning(const Learning&) = delete;
  Learning& operator=(const Learning&) = delete;

  // learning.cpp:69-90: image of explicit unit-box points / normals (convertToImageRGB with aligns = false).
  // The grasp argument only carries the binormal into the Instance, which the image does not use; idx > -1
  // (write a jpeg to a fixed path) and is_plotting are left out.
  ag2shim::Image createGraspImage(const ag2shim::Matrix3Xd& points, const ag2shim::Matrix3Xd& normals,
                                  const GraspHypothesis& /*grasp*/, int /*idx*/ = -1, bool /*is_plotting*/ = false) {
    if (points.cols() != normals.cols()) throw std::runtime_error("Learning: points / normals size mismatch");
    if (!own_) {
      ag2_params p;
      ag2_default_params(&p);
      ag2shim::check(ag2_create(&p, 0, &own_), "ag2_create");
    }
    const uint32_t off[2] = {0u, (uint32_t)points.cols()};
    ag2shim::Image img;
    ag2shim::check(ag2_render_point_sets(own_, points.data(), normals.data(), off, 1, 0, img.data.data()),
                   "ag2_render_point_sets", own_);
    return img;
  }

  // learning.cpp:4-33.  The hands must come from the last generateHypotheses / detect call of ONE
  // context (they refer to its in-box points); cam_pos and the plot flags are unused, as in the reference.
  std::vector<ag2shim::Image> createGraspImages(const std::vector<GraspHypothesis>& hands_list,
                                                const ag2shim::Matrix3Xd& /*cam_pos*/, bool /*is_plotting*/ = false,
                                                bool /*is_plotting_best*/ = false) {
    std::vector<ag2shim::Image> out(hands_list.size());
    if (hands_list.empty()) return out;
    ag2_ctx* ctx = const_cast<ag2_ctx*>(hands_list[0].origin());
    if (!ctx) throw std::runtime_error("Learning: hypotheses without an originating context");
    std::vector<int32_t> sid(hands_list.size()), oid(hands_list.size());
    for (size_t i = 0; i < hands_list.size(); i++) {
      if (hands_list[i].origin() != ctx) throw std::runtime_error("Learning: hypotheses of different contexts");
      sid[i] = hands_list[i].record().sample_id;
      oid[i] = hands_list[i].record().orientation_id;
    }
    std::vector<uint8_t> buf(hands_list.size() * (size_t)AG2_IMAGE_BYTES);
    ag2shim::check(ag2_render_hypotheses(ctx, sid.data(), oid.data(), sid.size(), buf.data()),
                   "ag2_render_hypotheses", ctx);
    for (size_t i = 0; i < out.size(); i++)
      out[i].data.assign(buf.begin() + i * AG2_IMAGE_BYTES, buf.begin() + (i + 1) * AG2_IMAGE_BYTES);
    return out;
  }

 private:
  int size_, num_threads_;
  ag2_ctx* own_;  // context of the point-set path (the hypothesis path uses the context the hands came from)
};

typedef std::pair<std::string, float> Prediction;  // caffe_classifier.h:52

class Classifier {
 public:
  // caffe_classifier.cpp:4-35.  model_file (prototxt) is fixed to the shipped LeNet; trained_file is
  // the caffemodel; an empty trained_file draws the prototxt's Xavier fillers (seed 1) instead.
  Classifier(const std::string& /*model_file*/, const std::string& trained_file, const std::string& /*label_file*/)
      : ctx_(nullptr) {
    ag2_params p;
    ag2_default_params(&p);
    ag2shim::check(ag2_create(&p, 0, &ctx_), "ag2_create");
    int rc = trained_file.empty() ? ag2_init_lenet_xavier(ctx_, 1) : ag2_load_lenet_caffemodel(ctx_, trained_file.c_str());
    if (rc != AG2_OK) {
      std::string msg = std::string("Classifier: cannot load ") + trained_file + ": " + ag2_last_error(ctx_);
      ag2_destroy(ctx_);
      throw std::runtime_error(msg);  // the reference aborts through glog CHECK here
    }
    labels_.push_back("0");  // caffe/labels.txt
    labels_.push_back("1");
  }
  ~Classifier() { ag2_destroy(ctx_); }
  Classifier(const Classifier&) = delete;
  Classifier& operator=(const Classifier&) = delete;

  // caffe_classifier.cpp:70-91: per image, (label, ip2 logit) for each class
  std::vector<std::vector<Prediction> > ClassifyBatch(const std::vector<ag2shim::Image>& imgs, int num_classes) {
    std::vector<uint8_t> buf(imgs.size() * (size_t)AG2_IMAGE_BYTES);
    for (size_t i = 0; i < imgs.size(); i++) std::copy(imgs[i].data.begin(), imgs[i].data.end(), buf.begin() + i * AG2_IMAGE_BYTES);
    std::vector<float> logits(2 * imgs.size() + 2);
    ag2shim::check(ag2_classify_images(ctx_, buf.data(), imgs.size(), logits.data()), "ag2_classify_images", ctx_);
    std::vector<std::vector<Prediction> > out(imgs.size());
    for (size_t i = 0; i < imgs.size(); i++)
      for (int k = 0; k < num_classes && k < 2; k++) out[i].push_back(std::make_pair(labels_[k], logits[2 * i + k]));
    return out;
  }

  // caffe_classifier.cpp:57-67 (use_softmax is ignored there as well: the ip2 blob is read)
  std::vector<Prediction> Classify(const ag2shim::Image& img, bool /*use_softmax*/ = false) {
    return ClassifyBatch(std::vector<ag2shim::Image>(1, img), 2)[0];
  }

 private:
  ag2_ctx* ctx_;
  std::vector<std::string> labels_;
};

#endif
