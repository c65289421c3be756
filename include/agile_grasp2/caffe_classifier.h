// The reference keeps Classifier in caffe_classifier.h; the shim lives beside Learning.
#ifndef AGILE_GRASP2_SHIM_CAFFE_CLASSIFIER_H_
#define AGILE_GRASP2_SHIM_CAFFE_CLASSIFIER_H_
#include "learning.h"
#endif
