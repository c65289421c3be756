/*
 * ag2.h — C ABI of the B200-native grasp-hypothesis hot path (libag2.so).
 *
 * This is the drop-in boundary for ONE path of gwding/agile_grasp2: everything
 * GraspDetector::detectGraspPoses does between receiving a CloudCamera and
 * returning scored GraspHypothesis records (reference grasp_detector.cpp:84-282):
 *   neighbour search -> normals -> local frames -> hand search -> antipodal test
 *   -> prune -> grasp images -> LeNet -> score/select.
 *
 * The reference has no FFI of its own; its boundary is a C++ class surface
 * (hand_search.h:93-133, learning.h:75-81, caffe_classifier.h:55-66,
 * grasp_detector.h:72-126).  Every entry point below names the reference
 * interface it replaces.  include/agile_grasp2/ holds header-only C++ shims with the
 * reference's class names that forward to these functions.
 *
 * Conventions: extern "C", plain pointers and sizes, no torch/Eigen/PCL types.
 * Every function returns an int status: AG2_OK (0) or a negative AG2_E_* code;
 * ag2_strerror() names it.  One ctx == one GPU == one in-flight call; calls on a
 * ctx must be serialised by the caller (the reference's HandSearch is not
 * re-entrant either: it keeps cloud_normals_ as member state, hand_search.h:172).
 * The caller owns every host buffer it passes in or receives; the ctx owns all
 * device memory, streams and pinned staging.  There is NO CPU fallback: if no
 * CUDA device is usable ag2_create fails with AG2_E_NO_DEVICE.
 */
#ifndef AG2_H_
#define AG2_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AG2_ABI_VERSION 1

/* status codes */
enum {
  AG2_OK = 0,
  AG2_E_INVALID = -1,      /* bad argument / bad parameter value */
  AG2_E_NO_DEVICE = -2,    /* no usable CUDA device (no CPU fallback exists) */
  AG2_E_CUDA = -3,         /* a CUDA runtime call failed; see ag2_last_error */
  AG2_E_EMPTY_CLOUD = -4,  /* reference: empty vector (grasp_detector.cpp:86-91) */
  AG2_E_NO_SAMPLES = -5,
  AG2_E_NO_WEIGHTS = -6,   /* antipodal_mode 1 without classifier weights */
  AG2_E_STATE = -7,        /* stage called out of order */
  AG2_E_IO = -8,           /* file could not be read / parsed */
  AG2_E_CAPACITY = -9,     /* caller buffer too small */
  AG2_E_NOMEM = -10
};

/* antipodal_mode, reference grasp_detector.cpp:5-7 */
enum { AG2_MODE_NONE = 0, AG2_MODE_PREDICTION = 1, AG2_MODE_GEOMETRIC = 2 };

#define AG2_MAX_ORIENTATIONS 64
#define AG2_NUM_FINGER_SLOTS 20 /* 2 x 10 placements, finger_hand.cpp:7-13 */
#define AG2_MAX_DEEPEN_STEPS 62
#define AG2_IMAGE_SIZE 60       /* Learning(60, ...), grasp_detector.cpp:56 */
#define AG2_IMAGE_BYTES (60 * 60 * 3)

/*
 * Parameters.  Mirrors HandSearch::Parameters (hand_search.h:72-91) field for
 * field, followed by the GraspDetector members that steer the path
 * (grasp_detector.cpp:17-80).  ag2_default_params() fills the compiled-in
 * defaults of the reference.
 */
typedef struct ag2_params {
  /* HandSearch::Parameters */
  double nn_radius_taubin;    /* radius of the local-frame neighbourhood (0.01) */
  double nn_radius_hands;     /* radius of the hand-search neighbourhood (0.1) */
  int32_t num_orientations;   /* 8 */
  int32_t num_threads;        /* accepted; used only by CPU baselines */
  int32_t num_samples;        /* informational, as in the reference */
  int32_t min_inliers;        /* HandleSearch::min_inliers_ (grasp_detector.cpp:59-65); 0 = no clustering */
  double finger_width;        /* 0.01 */
  double hand_outer_diameter; /* 0.09 */
  double hand_depth;          /* 0.06 */
  double hand_height;         /* 0.02 */
  double init_bite;           /* 0.015 (launch files use 0.01) */
  double cam_tf_left[16];     /* 4x4 row-major camera pose */
  double cam_tf_right[16];
  /* GraspDetector */
  double workspace[6];        /* minX maxX minY maxY minZ maxZ */
  int32_t antipodal_mode;     /* AG2_MODE_* (reference default PREDICTION) */
  int32_t filter_half_grasps; /* bool, default 1 */
  int32_t prune;              /* bool: run S13; the reference skips it when the
                                 sample_indices param is set (grasp_detector.cpp:150-160) */
  int32_t num_selected;       /* 50; <=0 keeps all */
  double min_score_diff;      /* 500.0 */
  double min_aperture;        /* 0.03 */
  double max_aperture;        /* 0.07 */
  /* constants the reference hard-codes, exposed for the scaling studies */
  double normals_radius;      /* 0.01, hand_search.cpp:91 */
  uint64_t rng_seed;          /* counter-based replacement of rand() (hand_search.cpp:130) */
} ag2_params;

/*
 * One grasp hypothesis: the GraspMsg fields (msg/GraspMsg.msg:5-16,
 * grasp_hypothesis.cpp:37-50) in double precision plus provenance.  192 bytes.
 */
typedef struct ag2_hypothesis {
  double surface[3];
  double bottom[3];
  double top[3];
  double axis[3];
  double approach[3];
  double binormal[3];
  double width;
  double score;
  int32_t sample_id;       /* position in the sample list given to the ctx */
  int32_t orientation_id;  /* 0 .. num_orientations-1 */
  uint8_t half_antipodal;
  uint8_t full_antipodal;
  uint8_t finger_slot;     /* chosen placement e (0..9) */
  uint8_t depth_step;      /* deepen steps taken beyond init_bite */
  int32_t num_in_box;      /* points in the closing region */
  double hand_top;         /* FingerHand::top_    (hand frame, m) */
  double hand_bottom;      /* FingerHand::bottom_ (hand frame, m) */
} ag2_hypothesis;

/* Per (sample, orientation) trace of the finger sweep, for parity tests. */
typedef struct ag2_sweep_record {
  uint32_t fingers;     /* bit i: slot i collision-free at init_bite (finger_hand.cpp:17-93) */
  uint16_t hand;        /* bit i: fingers[i] && fingers[10+i] (finger_hand.cpp:313-325) */
  int8_t slot;          /* chosen e or -1 */
  int8_t depth_step;    /* last successful deepen step (0 = init_bite) or -1 */
  int32_t num_in_box;   /* -1 if the orientation was infeasible */
  uint8_t antipodal;    /* 0 NO, 1 HALF, 2 FULL (antipodal.cpp:3-5) */
  uint8_t emitted;      /* 1 if a hypothesis was emitted */
  uint16_t reserved;
} ag2_sweep_record;

/* Local frame of one sample (local_frame.h:53-146). */
typedef struct ag2_frame {
  double sample[3];
  double normal[3];
  double binormal[3];
  double curvature_axis[3];
  int32_t valid;           /* 0: no neighbours within nn_radius_taubin -> sample dropped */
  int32_t majority_cam;
  int32_t num_neighbors;   /* neighbours within nn_radius_taubin */
  int32_t reserved;
} ag2_frame;

/* Stage counters (the reference prints these, grasp_detector.cpp:142,155,222,253). */
typedef struct ag2_stats {
  int64_t num_points;
  int64_t num_samples;
  int64_t num_frames;
  int64_t num_hypotheses;       /* emitted by the hand search */
  int64_t num_pruned;           /* survivors of S13 */
  int64_t num_antipodal;        /* survivors of mode 1 / 2 */
  int64_t num_selected;
  int64_t sum_neighbors_hands;  /* sum over samples of K */
  int64_t sum_cropped;          /* sum over samples of k' */
  int64_t sum_in_box;           /* sum over hypotheses of m */
  int64_t sum_neighbors_normals;/* sum over points of k_n */
  int64_t gpu_launches;         /* kernels launched since the last ag2_set_cloud */
  double ms_phase[8];           /* grid, normals, frames, hands, prune, render, lenet, select */
  double ms_gap[8];             /* stream time between the end of the previous timed phase and the start of this
                                   one (host round trips, copies, untimed launches) */
} ag2_stats;

typedef struct ag2_ctx ag2_ctx;

/* ---- lifecycle ---------------------------------------------------------- */

/* Fills the reference's compiled-in defaults (grasp_detector.cpp:17-80, the
 * Baxter camera poses of :111-127, workspace of launch/file_detect_grasps.launch:16). */
int ag2_default_params(ag2_params* p);

/* Replaces: GraspDetector ctor + HandSearch::setParameters (grasp_detector.cpp:15-81,
 * hand_search.cpp:64-80).  device = CUDA ordinal. */
int ag2_create(const ag2_params* p, int device, ag2_ctx** out);
/* ~GraspDetector (grasp_detector.h:76-80): frees the device memory, streams and events of the ctx. */
int ag2_destroy(ag2_ctx* ctx);
/* HandSearch::setParameters / setCamTfLeft / setCamTfRight (hand_search.h:121-133). */
int ag2_set_params(ag2_ctx* ctx, const ag2_params* p);
/* Error reporting of the C ABI (no reference counterpart: the reference prints and returns an empty list, e.g.
 * grasp_detector.cpp:86-91): every entry point returns an AG2_E_* code; text for the code and for the last failure. */
const char* ag2_strerror(int code);
const char* ag2_last_error(const ag2_ctx* ctx);
/* AG2_ABI_VERSION the library was built with (no reference counterpart). */
int ag2_abi_version(void);

/* ---- classifier weights ------------------------------------------------- */

/* Replaces Classifier ctor's net_->CopyTrainedLayersFrom (caffe_classifier.cpp:13-14):
 * parses the caffemodel protobuf wire format (V1 "layers" and V2 "layer"). */
int ag2_load_lenet_caffemodel(ag2_ctx* ctx, const char* path);
/* The parser behind it on its own (host code, no ctx, no GPU): the eight blobs of a .caffemodel into caller
 * arrays in the layouts of ag2_set_lenet_weights (conv1 20x1500/20 ... see below; NULL = not wanted).  Layers are
 * matched by the names of caffe/sigmoid_lenet.prototxt:10,34,58,76 (conv1, conv2, ip1, ip2); a layer without a
 * bias blob gets zeros; AG2_E_IO for an unreadable or malformed file, a missing layer or blob sizes that are not
 * the prototxt's. */
int ag2_read_lenet_caffemodel(const char* path, float* conv1_w, float* conv1_b, float* conv2_w, float* conv2_b,
                              float* ip1_w, float* ip1_b, float* ip2_w, float* ip2_b);
/* The same blobs from memory (what net_->CopyTrainedLayersFrom leaves in the four layers of
 * caffe/sigmoid_lenet.prototxt:10-90).  Caffe blob layouts: conv [out,in,kh,kw], ip [out,in]; fp32. */
int ag2_set_lenet_weights(ag2_ctx* ctx, const float* conv1_w, const float* conv1_b,
                          const float* conv2_w, const float* conv2_b, const float* ip1_w,
                          const float* ip1_b, const float* ip2_w, const float* ip2_b);
/* Xavier fillers named in caffe/sigmoid_lenet.prototxt:17-19,41-43,64-66,82-84
 * (the trained blob is absent from the reference checkout). */
int ag2_init_lenet_xavier(ag2_ctx* ctx, uint64_t seed);
/* Reads the blobs back in the same layouts (tests; no reference counterpart). */
int ag2_get_lenet_weights(ag2_ctx* ctx, float* conv1_w, float* conv1_b, float* conv2_w,
                          float* conv2_b, float* ip1_w, float* ip1_b, float* ip2_w,
                          float* ip2_b);

/* ---- inputs (the CloudCamera contract, cloud_camera.h:102-139) ----------- */

/* Processed cloud.  xyz: n points, stride_bytes apart (12 for packed floats, 32
 * for pcl::PointXYZRGBA).  cam_mask[i] bit c = point i seen by camera c
 * (camera_source_, cloud_camera.cpp:39-50); num_cams = rows of camera_source_.
 * normals_or_null: optional precomputed normals, 3 floats per point
 * (CloudCamera::getNormals, hand_search.cpp:20-29). Host pointers.  The copies are asynchronous:
 * PINNED host buffers must stay valid and unchanged until the next call on this ctx that returns
 * results (any stage call synchronises); pageable buffers may be reused at once.
 * Call order: the cloud first, then the samples -- a new cloud drops sample INDICES set before
 * (they were validated against the old cloud; ag2_detect then returns AG2_E_NO_SAMPLES). */
int ag2_set_cloud(ag2_ctx* ctx, const float* xyz, size_t stride_bytes, size_t n,
                  const uint8_t* cam_mask, int num_cams, const float* normals_or_null);
/* Same contract (cloud_camera.h:102-139) with DEVICE pointers (packed xyz; used after an NCCL broadcast). */
int ag2_set_cloud_device(ag2_ctx* ctx, const float* d_xyz, size_t n, const uint8_t* d_cam_mask,
                         int num_cams, const float* d_normals_or_null);
/* CloudCamera::setSampleIndices (cloud_camera.h:132-135) -> calculateLocalFrames(indices),
 * hand_search.cpp:97-170. */
int ag2_set_sample_indices(ag2_ctx* ctx, const int32_t* idx, size_t n);
/* CloudCamera::setSampleIndices (cloud_camera.h:132-135) with a DEVICE pointer (the sample list after an NCCL
 * broadcast); indices are range-checked on the device. */
int ag2_set_sample_indices_device(ag2_ctx* ctx, const int32_t* d_idx, size_t n);
/* CloudCamera::setSamples (cloud_camera.cpp:254-257) -> calculateLocalFrames(samples),
 * hand_search.cpp:238-317.  xyz: 3 doubles per sample. */
int ag2_set_sample_points(ag2_ctx* ctx, const double* xyz, size_t n);
/* Offset added to the sample position when deriving the stream that replaces rand() in
 * hand_search.cpp:130, so that a shard of a larger sample list draws like the whole list
 * (no reference counterpart: its rand() is shared by the OpenMP threads). */
int ag2_set_sample_id_offset(ag2_ctx* ctx, int64_t first_sample_id);
/* General form (no reference counterpart either): sample s of this context is sample first + stride*s of the
 * whole list (a strided shard; stride >= 1).  ag2_set_sample_id_offset(first) == ag2_set_sample_id_map(first, 1). */
int ag2_set_sample_id_map(ag2_ctx* ctx, int64_t first_sample_id, int64_t stride);

/* ---- multi-GPU: the normals (calculateNormalsOMP, hand_search.cpp:83-94) sharded by point ------
 * Every rank holds the same cloud.  Rank r of w computes the normals of the 32-point granules
 * g % w == r of the cell-sorted cloud into caller-owned DEVICE memory (4 floats per point,
 * cap_points >= *shard_points), the ranks all-gather those buffers (NCCL), and
 * ag2_set_normals_shards installs the gathered w x cap_points x 4 floats.  Afterwards the
 * context is exactly as after ag2_compute_normals. */
int ag2_normals_shard_points(ag2_ctx* ctx, int world, size_t* shard_points /* of rank 0 = the largest */);
int ag2_compute_normals_shard(ag2_ctx* ctx, int rank, int world, float* d_out, size_t cap_points, size_t* n_owned);
int ag2_set_normals_shards(ag2_ctx* ctx, const float* d_all, size_t cap_points, int world);

/* ---- pre-processing (next row f1) ----------------------------------------- */

/* GraspDetector::preprocessPointCloud steps 1-2 (grasp_detector.cpp:289-303):
 * CloudCamera::filterWorkspace (cloud_camera.cpp:89-121, bounds = params.workspace) and
 * CloudCamera::voxelizeCloud (cloud_camera.cpp:124-168, voxel_size 0.003 in the reference) on
 * the ORIGINAL cloud given as host buffers; the processed cloud becomes the ctx's cloud exactly
 * as if ag2_set_cloud had been called with it.  Precomputed normals follow the workspace filter
 * and are dropped by voxelisation (the reference leaves them stale). */
int ag2_preprocess_cloud(ag2_ctx* ctx, const float* xyz, size_t stride_bytes, size_t n,
                         const uint8_t* cam_mask, int num_cams, const float* normals_or_null,
                         int filter_workspace, int voxelize, double voxel_size);
/* The ctx's (processed) cloud: packed xyz and camera masks.  CloudCamera::getCloudProcessed /
 * getCameraSource (cloud_camera.h:102-110). */
int ag2_get_cloud(ag2_ctx* ctx, float* xyz, uint8_t* cam_mask, size_t cap, size_t* n);
/* CloudCamera::subsampleUniformly (cloud_camera.cpp:171-178): num_samples distinct ascending
 * indices into the processed cloud become the sample indices. */
int ag2_subsample_uniformly(ag2_ctx* ctx, int num_samples, uint64_t seed);
/* CloudCamera::getSampleIndices (cloud_camera.h:117-120). */
int ag2_get_sample_indices(ag2_ctx* ctx, int32_t* idx, size_t cap, size_t* n);

/* ---- stages -------------------------------------------------------------- */

/* KdTreeFLANN::setInputCloud + calculateNormalsOMP (hand_search.cpp:11-30,83-94). */
int ag2_compute_normals(ag2_ctx* ctx);
/* HandSearch::generateHypotheses (hand_search.cpp:4-61): frames + evaluateHands. */
int ag2_generate_hypotheses(ag2_ctx* ctx);
/* GraspDetector::pruneGraspsOnHandParameters (grasp_detector.cpp:363-395). */
int ag2_prune(ag2_ctx* ctx);
/* Learning::createGraspImages (learning.cpp:4-33) on the pruned list.
 * out_or_null: host buffer of num_pruned * AG2_IMAGE_BYTES (HWC u8, the cv::Mat layout). */
int ag2_render_images(ag2_ctx* ctx, uint8_t* out_or_null);
/* Learning::createGraspImages (learning.cpp:4-33) for caller-chosen hypotheses of the last ag2_generate_hypotheses call, addressed by
 * (sample_id, orientation_id) as found in their records; out: n * AG2_IMAGE_BYTES (host). */
int ag2_render_hypotheses(ag2_ctx* ctx, const int32_t* sample_ids, const int32_t* orientation_ids,
                          size_t n, uint8_t* out);
/* Learning::createGraspImage(points, normals, grasp) (learning.cpp:69-90): images of caller-supplied
 * unit-box point sets, the generator behind the ClassificationNode service
 * (classification_node.cpp:75-141).  Image i is drawn from points / normals
 * [offsets[i], offsets[i+1]) (3 doubles each, host pointers); aligns = 0 reproduces the reference
 * (convertToImageRGB(ins, false)), aligns != 0 subtracts the minimum y first as the detection
 * path does.  out: n_images x 10800 bytes (60 x 60 x 3, HWC). */
int ag2_render_point_sets(ag2_ctx* ctx, const double* pts, const double* normals, const uint32_t* offsets,
                          size_t n_images, int aligns, uint8_t* out);
/* Classifier::ClassifyBatch + thresholding (grasp_detector.cpp:186-209). */
int ag2_classify(ag2_ctx* ctx);
/* mode 2: keep isFullAntipodal (grasp_detector.cpp:212-221); mode 0: keep all. */
int ag2_select(ag2_ctx* ctx);
/* HandleSearch::findClusters (handle_search.cpp:4-80, remove_inliers = false) on list
 * `which_src`; result in AG2_LIST_CLUSTERS (records in source order).  ag2_select runs it on the
 * antipodal list before the top-k when params.min_inliers > 0 (grasp_detector.cpp:229-236). */
int ag2_find_clusters(ag2_ctx* ctx, int which_src);
/* HandleSearch::setMinInliers (handle_search.h:83-86): the cluster size threshold alone, without the invalidation of
 * the lists that ag2_set_params implies; used by the next ag2_find_clusters / ag2_detect. */
int ag2_set_min_inliers(ag2_ctx* ctx, int min_inliers);
/* GraspDetector::detectGraspPoses (grasp_detector.cpp:84-282), clustering included when
 * params.min_inliers > 0. */
int ag2_detect(ag2_ctx* ctx);

/* ImportanceSampling (importance_sampling.cpp:9-15,30-112).  Constants of the reference are the
 * defaults of ag2_default_is_params; c4 scales num_samples. */
typedef struct ag2_is_params {
  int32_t num_iterations;     /* 5 */
  int32_t num_samples;        /* 50 per round */
  int32_t num_init_samples;   /* 100 (unused by the reference as well) */
  int32_t method;             /* 1 = sum of Gaussians, 2 = max of Gaussians (default) */
  double prob_rand_samples;   /* 0.3 */
  double radius;              /* 0.02: sigma of the Gaussians */
  uint64_t seed;              /* counter-based replacement of rand() / boost::mt19937(time) */
} ag2_is_params;
/* The constants of importance_sampling.cpp:9-15. */
int ag2_default_is_params(ag2_is_params* p);
/* ImportanceSampling::detectGraspPoses (importance_sampling.cpp:30-112) without the final
 * clustering: detect on the samples currently set, then num_iterations rounds of
 * {draw samples near the grasps found so far + random cloud points, detect, append}.
 * The grid and the normals are built once and reused by every round.  Result:
 * AG2_LIST_IMPORTANCE (records of all rounds; sample_id counts through the rounds). */
int ag2_detect_importance(ag2_ctx* ctx, const ag2_is_params* p);
/* One step of the above for callers that run the rounds themselves (the multi-GPU host shards every
 * round's samples over the ranks and merges the selections): draws the num_samples sample points of
 * round `round` (0-based) from `hands` = all hands found so far (host records, n_hands > 0), exactly as
 * ag2_detect_importance does (importance_sampling.cpp:47-86, :115-157); xyz_out: 3 * num_samples
 * doubles on the host.  Needs the cloud (random part); does not touch the ctx's samples or lists. */
int ag2_draw_importance_samples(ag2_ctx* ctx, const ag2_is_params* p, int round, const ag2_hypothesis* hands,
                                size_t n_hands, double* xyz_out);
/* Samples drawn in round `round` (0-based), 3 doubles each, for the parity tests. */
int ag2_get_importance_samples(ag2_ctx* ctx, int round, double* xyz, size_t cap, size_t* n);
/* Waits for everything enqueued on the ctx stream (no reference counterpart: the reference is synchronous). */
int ag2_synchronize(ag2_ctx* ctx);
/* The CUDA stream (cudaStream_t) every kernel of this ctx is launched on, for event timing (no reference
 * counterpart). */
int ag2_get_stream(ag2_ctx* ctx, void** stream);

/* Measurement aid (no reference counterpart): FMA throughput of this ctx's GPU right now, in TFLOP/s --
 * the denominators of the issue-bound rooflines bench.py reports for the normals and hand-sweep kernels
 * ("FP64 peak from a same-run FMA microbenchmark"). */
int ag2_measure_fp_peaks(ag2_ctx* ctx, double* fp32_tflops, double* fp64_tflops);

/* ---- outputs ------------------------------------------------------------- */

enum {
  AG2_LIST_HYPOTHESES = 0,
  AG2_LIST_PRUNED = 1,
  AG2_LIST_ANTIPODAL = 2,
  AG2_LIST_SELECTED = 3,
  AG2_LIST_IMPORTANCE = 4, /* result of ag2_detect_importance */
  AG2_LIST_CLUSTERS = 5    /* result of ag2_find_clusters */
};
/* The std::vector<GraspHypothesis> each step of GraspDetector::detectGraspPoses holds (grasp_detector.cpp:
 * 141 hands, :149-160 hands_filtered, :170-221 antipodal_hands, :239-252 hands_selected), as ag2_hypothesis
 * records (grasp_hypothesis.h:62-313) in the reference's list order. */
int ag2_num_hypotheses(ag2_ctx* ctx, int which, size_t* n);
int ag2_get_hypotheses(ag2_ctx* ctx, int which, ag2_hypothesis* out, size_t cap, size_t* n);
/* Device pointer of a list (records are contiguous); valid until the next stage call (no reference counterpart). */
int ag2_get_hypotheses_device(ag2_ctx* ctx, int which, const ag2_hypothesis** d_ptr, size_t* n);
/* Copies a list into caller-owned DEVICE memory of `cap` records and zero-fills the rest (a valid
 * record never has an all-zero approach vector); the buffer is complete when the call returns.
 * This is the send buffer of the multi-GPU all-gather (grasp lists of all ranks; no reference counterpart). */
int ag2_copy_hypotheses_device(ag2_ctx* ctx, int which, ag2_hypothesis* d_dst, size_t cap, size_t* n);
/* Counts and per-phase device times of the last call: what the reference prints with omp_get_wtime
 * (grasp_detector.cpp:93,174,185,210,254; hand_search.cpp:7,19,30,58,166,230). */
int ag2_get_stats(ag2_ctx* ctx, ag2_stats* out);

/* ---- introspection used by the parity tests (intermediate values the reference keeps in members / locals) -- */
/* HandSearch::cloud_normals_ (hand_search.cpp:93), as floats. */
int ag2_get_normals(ag2_ctx* ctx, float* out /* 3*n, input point order */);
/* The LocalFrame of every sample (local_frame.h:53-146; hand_search.cpp:150-152), invalid ones included. */
int ag2_get_frames(ag2_ctx* ctx, ag2_frame* out, size_t cap);
/* FingerHand's state after calculateHand for every (sample, orientation) (finger_hand.h:201-222; hand_search.cpp:322,363-376). */
int ag2_get_sweep_records(ag2_ctx* ctx, ag2_sweep_record* out, size_t cap /* samples*R */);
/* Fixed-radius query (KdTreeFLANN::radiusSearch as called in hand_search.cpp:122,201,268): ids in
 * input point order sorted by (d2, id).  Host pointers. */
int ag2_radius_search(ag2_ctx* ctx, const float* q /*3*/, double radius, int32_t* idx, float* d2,
                      size_t cap, size_t* n);
/* In-box point ids (input order, reference list order) of hypothesis h of list `which`:
 * FingerHand::computePointsInClosingRegion (finger_hand.cpp:137-180) -> GraspHypothesis::indices_points_for_learning_. */
int ag2_get_in_box(ag2_ctx* ctx, int which, size_t h, int32_t* idx, size_t cap, size_t* n);
/* ip2 logits of the pruned list after ag2_classify: 2 floats per image (the `predictions` of grasp_detector.cpp:196). */
int ag2_get_logits(ag2_ctx* ctx, float* out, size_t cap_images);
/* Classifier::ClassifyBatch (caffe_classifier.cpp:70-127) on caller images (HWC u8, 60x60x3): 2 floats per image. */
int ag2_classify_images(ag2_ctx* ctx, const uint8_t* images, size_t n, float* logits_out);
/* The same classifier pass for callers that may not allocate (CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * d_images (n x 10800 bytes, HWC u8) and d_logits (2 floats per image) are
 * DEVICE buffers of the ctx's GPU, owned by the caller; the activations live in a fixed array inside
 * libag2.so (about 11.6 MB, held by every GPU that loads the library until exit), so the call makes
 * no cudaMalloc / cudaFree and at most AG2_CLASSIFY_STATIC_MAX_IMAGES images go in one call
 * (more: AG2_E_CAPACITY, nothing launched).  The kernels are enqueued on the ctx stream and the call
 * returns without waiting: the logits are ready after ag2_synchronize, or for work ordered behind
 * ag2_get_stream's stream.  The weights must have been set before (that step allocates).  All calls
 * on one GPU, from any ctx, share the array: two calls that overlap on the GPU (two ctx of one
 * device) race; calls on one ctx are ordered by its stream.  Logits equal ag2_classify_images'. */
#define AG2_CLASSIFY_STATIC_MAX_IMAGES 64
int ag2_classify_images_static(ag2_ctx* ctx, const uint8_t* d_images, size_t n, float* d_logits);

/* ---- on-disk / wire formats (next row f3; host code, no ctx needed) -------- */

/* pcl::io::loadPCDFile (cloud_camera.cpp:231-240): DATA ascii | binary | binary_compressed;
 * fields x y z [rgb|rgba] [normal_x normal_y normal_z]; other fields are skipped. */
int ag2_pcd_info(const char* path, size_t* n, int* has_normals, int* has_color);
int ag2_pcd_read(const char* path, float* xyz /*3*cap*/, float* normals_or_null, uint32_t* rgba_or_null,
                 size_t cap, size_t* n);
/* pcl::io::savePCDFile's format for the tests and tools (no reference counterpart: it only reads,
 * cloud_camera.cpp:231-240). */
int ag2_pcd_write(const char* path, const float* xyz, const float* normals_or_null,
                  const uint32_t* rgba_or_null, size_t n, int binary);
/* ROS 1 wire format of agile_grasp2/GraspListMsg (msg/GraspListMsg.msg, msg/GraspMsg.msg;
 * GraspHypothesis::convertToGraspMsg, grasp_hypothesis.cpp:37-50).  out may be NULL to size. */
int ag2_serialize_grasp_list(const ag2_hypothesis* hands, size_t n, uint32_t seq, uint32_t stamp_sec,
                             uint32_t stamp_nsec, const char* frame_id, uint8_t* out, size_t cap, size_t* len);

/* Multi-GPU host: merges the per-rank SELECTED blocks of an all-gather (`world` blocks of `cap` records,
 * block r from the rank that evaluated samples r, r + world, ...; records with an all-zero approach
 * vector are padding) into the single-GPU selection: sample_id -> r + world * sample_id, list order
 * (sample, orientation), then the top num_selected by score with ties in list order
 * (grasp_detector.cpp:238-252; num_selected <= 0 keeps all).  Host-only, needs no ctx.  out: world * cap records. */
int ag2_merge_selected(const ag2_hypothesis* blocks, int world, size_t cap, int num_selected, ag2_hypothesis* out,
                       size_t* n_out);


/* ---- multi-GPU: the same path over 1..8 GPUs of one box --------------------------
 * The reference's caller hands ONE cloud to GraspDetector::detectGraspPoses /
 * ImportanceSampling::detectGraspPoses (grasp_detection_node.cpp:109-115,135-138) and gets one list
 * back; its only parallelism is OpenMP over samples (hand_search.cpp:118,194).  A group keeps that
 * contract over several GPUs: rank 0 supplies the cloud and the sample list, every rank returns the
 * same merged selection, byte for byte what one GPU returns for the whole list.  Inside, the path
 * is sharded by grasp sample (strided shards), the normals by 32-point granules of the cell-sorted
 * cloud, and the exchanges run on peer memory over NVLink: ranks pull rank 0's staged cloud, the
 * normals kernel stores each result into the arrays of all ranks, the selected records are stored
 * into every rank's result window; device-side flags order the steps, no host synchronisation
 * between ranks.  NCCL bootstraps the group and is the fallback transport (AG2_GROUP_NCCL=1).
 * One rank = one GPU = one ag2_ctx (classifier weights are set on the ctx of every LOCAL rank). */
typedef struct ag2_group ag2_group;
typedef struct ag2_group_id { unsigned char bytes[128]; } ag2_group_id;  /* an ncclUniqueId */

/* One process, n_gpus ranks on devices 0..n_gpus-1, one host thread per GPU inside the library: the place of
 * the GraspDetector the node constructs once (grasp_detection_node.cpp:98-107).  This is what the C++ shim's
 * GraspDetector uses when Params::num_gpus > 1. */
int ag2_group_create(const ag2_params* p, int n_gpus, ag2_group** out);
/* One process per rank (torchrun, mpirun; no reference counterpart): rank 0 calls ag2_group_unique_id and hands
 * the id to the others by any channel; then every rank calls ag2_group_create_rank (collective). */
int ag2_group_unique_id(ag2_group_id* out);
int ag2_group_create_rank(const ag2_params* p, const ag2_group_id* id, int rank, int world, int device,
                          ag2_group** out);
/* Lifetime and shape of the group (no reference counterpart). */
int ag2_group_destroy(ag2_group* g);
int ag2_group_size(const ag2_group* g, int* world, int* n_local);
/* ctx of the i-th LOCAL rank: weights (ag2_set_lenet_weights ...), ag2_get_stats, ag2_get_stream
 * (no reference counterpart). */
int ag2_group_local_ctx(ag2_group* g, int i, ag2_ctx** ctx);
const char* ag2_group_last_error(const ag2_group* g);
/* 1 when the exchanges run on peer memory (NVLink loads / stores), 0 on the NCCL fallback; valid after
 * the first detect call (measurement aid, no reference counterpart). */
int ag2_group_transport(const ag2_group* g, int* peer_stores);
/* Bytes the FIRST local rank moved between host and device in the last call (cloud + samples in, result
 * window out); every other rank of a multi-process group reports h2d = 0 (measurement aid, no reference
 * counterpart). */
int ag2_group_io_bytes(const ag2_group* g, size_t* h2d, size_t* d2h);
/* Per-segment account of the FIRST local rank's last ag2_group_detect call, ms[10]: device time (CUDA events on
 * the ctx stream) of [0] inputs staged and pulled, [1] cell index, [2] own normal moments, [3] wait for all cell
 * indexes + own normals stored to all ranks, [4] wait for all normals + sample shard, [5] local detect (frames ..
 * select), [6] results stored to all ranks + wait; host clock of [7] result exchange incl. D2H and merge, [8] the
 * whole call; [9] device time start .. results in (measurement aid, no reference counterpart). */
int ag2_group_last_timing(const ag2_group* g, double* ms);
/* GraspDetector::detectGraspPoses (grasp_detector.cpp:84-282) over the group.  Collective when ranks are
 * processes: every rank calls it with the same n, num_cams, n_samples; xyz (packed, 3 floats per point),
 * cam_mask (or NULL) and samples (indices into the cloud) are read on rank 0 only -- host pointers, or
 * device pointers on rank 0's GPU when on_device != 0.  out: the merged selection (top num_selected by
 * score, ties in (sample, orientation) order), identical on every rank; sample_id counts in the
 * caller's list. */
int ag2_group_detect(ag2_group* g, const float* xyz, size_t n, const uint8_t* cam_mask, int num_cams,
                     const int32_t* samples, size_t n_samples, int on_device, ag2_hypothesis* out, size_t cap,
                     size_t* n_out);
/* ImportanceSampling::detectGraspPoses (importance_sampling.cpp:30-112) over the group: the cloud, its
 * index and its normals stay resident on every rank; every round's samples are sharded like a plain
 * call and all ranks draw the same next batch from the merged hands.  out: the hands of all rounds
 * (ag2_detect_importance's AG2_LIST_IMPORTANCE), or their clusters when params.min_inliers > 0
 * (:105-109). */
int ag2_group_detect_importance(ag2_group* g, const float* xyz, size_t n, const uint8_t* cam_mask, int num_cams,
                                const int32_t* samples, size_t n_samples, int on_device, const ag2_is_params* ip,
                                ag2_hypothesis* out, size_t cap, size_t* n_out);

#ifdef __cplusplus
}
#endif
#endif /* AG2_H_ */
