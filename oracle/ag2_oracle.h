/*
 * ag2_oracle.h — C ABI of the CPU ORACLE (test infrastructure, NOT product code).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  It mirrors include/ag2.h call for call
 * (prefix ag2o_) so the parity tests drive both sides with the same inputs.
 * See ag2_oracle.cpp for the restatement and its reference citations.
 */
#ifndef AG2_ORACLE_H_
#define AG2_ORACLE_H_

#include "../include/ag2.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ag2o_ctx ag2o_ctx;

int ag2o_create(const ag2_params* p, ag2o_ctx** out);
int ag2o_destroy(ag2o_ctx* ctx);
int ag2o_set_params(ag2o_ctx* ctx, const ag2_params* p);
int ag2o_set_lenet_weights(ag2o_ctx* ctx, const float* conv1_w, const float* conv1_b,
                           const float* conv2_w, const float* conv2_b, const float* ip1_w,
                           const float* ip1_b, const float* ip2_w, const float* ip2_b);
int ag2o_set_cloud(ag2o_ctx* ctx, const float* xyz, size_t stride_bytes, size_t n,
                   const uint8_t* cam_mask, int num_cams, const float* normals_or_null);
int ag2o_set_sample_indices(ag2o_ctx* ctx, const int32_t* idx, size_t n);
int ag2o_set_sample_points(ag2o_ctx* ctx, const double* xyz, size_t n);
int ag2o_set_sample_id_offset(ag2o_ctx* ctx, int64_t first_sample_id);
int ag2o_set_sample_id_map(ag2o_ctx* ctx, int64_t first_sample_id, int64_t stride);
int ag2o_render_point_sets(ag2o_ctx* ctx, const double* pts, const double* normals, const uint32_t* offsets,
                           size_t n_images, int aligns, uint8_t* out);
int ag2o_point_normals(ag2o_ctx* ctx, const int32_t* idx, size_t n, float* out3);

int ag2o_compute_normals(ag2o_ctx* ctx);
int ag2o_generate_hypotheses(ag2o_ctx* ctx);
int ag2o_prune(ag2o_ctx* ctx);
int ag2o_render_images(ag2o_ctx* ctx, uint8_t* out_or_null);
int ag2o_classify(ag2o_ctx* ctx);
int ag2o_select(ag2o_ctx* ctx);
int ag2o_detect(ag2o_ctx* ctx);

int ag2o_preprocess_cloud(ag2o_ctx* ctx, const float* xyz, size_t stride_bytes, size_t n,
                          const uint8_t* cam_mask, int num_cams, const float* normals_or_null,
                          int filter_workspace, int voxelize, double voxel_size);
int ag2o_get_cloud(ag2o_ctx* ctx, float* xyz, uint8_t* cam_mask, size_t cap, size_t* n);
int ag2o_subsample_uniformly(ag2o_ctx* ctx, int num_samples, uint64_t seed);
int ag2o_get_sample_indices(ag2o_ctx* ctx, int32_t* idx, size_t cap, size_t* n);
int ag2o_find_clusters(ag2o_ctx* ctx, int which_src);
int ag2o_default_is_params(ag2_is_params* p);
int ag2o_detect_importance(ag2o_ctx* ctx, const ag2_is_params* p);
int ag2o_draw_importance_samples(ag2o_ctx* ctx, const ag2_is_params* p, int round, const ag2_hypothesis* hands,
                                 size_t n_hands, double* xyz_out);
int ag2o_get_importance_samples(ag2o_ctx* ctx, int round, double* xyz, size_t cap, size_t* n);

int ag2o_num_hypotheses(ag2o_ctx* ctx, int which, size_t* n);
int ag2o_get_hypotheses(ag2o_ctx* ctx, int which, ag2_hypothesis* out, size_t cap, size_t* n);
int ag2o_get_stats(ag2o_ctx* ctx, ag2_stats* out);
int ag2o_get_normals(ag2o_ctx* ctx, float* out);
int ag2o_get_frames(ag2o_ctx* ctx, ag2_frame* out, size_t cap);
int ag2o_get_sweep_records(ag2o_ctx* ctx, ag2_sweep_record* out, size_t cap);
/* brute != 0: O(N) scan (ground truth); 0: the bucketed kd-tree used by the timed baseline. */
int ag2o_radius_search(ag2o_ctx* ctx, const float* q, double radius, int brute, int32_t* idx,
                       float* d2, size_t cap, size_t* n);
int ag2o_get_in_box(ag2o_ctx* ctx, int which, size_t h, int32_t* idx, size_t cap, size_t* n);
int ag2o_get_logits(ag2o_ctx* ctx, float* out, size_t cap_images);
int ag2o_classify_images(ag2o_ctx* ctx, const uint8_t* images, size_t n, float* logits_out);

/* Stand-alone pieces exposed so pytest can pin them against real libraries here
 * (numpy.linalg.eigh, cv2.dilate / cvtColor / convertTo, torch conv2d). */
/* PCL eigen33 restatement: smallest eigenpair of a symmetric 3x3 float matrix. */
int ag2o_eigen33(const float* cov9, float* eigenvalue, float* eigenvector3);
/* LocalFrame::findAverageNormalAxis on n unit normals (3 doubles each). */
int ag2o_average_normal_axis(const double* normals, int n, const double* sample3,
                             const double* cam_origin3, double* normal3, double* binormal3,
                             double* axis3);
/* S14 tail: float HWC image -> dilate 3x3 -> BGR2RGB -> convertTo(u8, 255). */
int ag2o_image_postprocess(const float* img_hwc, uint8_t* out_hwc);
/* Host-side constant tables exactly as the reference builds them. */
int ag2o_finger_spacing(double finger_width, double hand_outer_diameter, double* out20);
int ag2o_deepen_depths(double init_bite, double hand_depth, double* out, int cap, int* n);
int ag2o_orientation_angles(int num_orientations, double* out);

#ifdef __cplusplus
}
#endif
#endif
