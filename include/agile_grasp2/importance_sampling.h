// The reference keeps ImportanceSampling in importance_sampling.h; the shim lives beside GraspDetector.
#ifndef AGILE_GRASP2_SHIM_IMPORTANCE_SAMPLING_H_
#define AGILE_GRASP2_SHIM_IMPORTANCE_SAMPLING_H_
#include "grasp_detector.h"
#endif
