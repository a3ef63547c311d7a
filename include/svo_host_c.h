/* svo_host_c.h - C entry points of libsvo_host.so: the C++ host layer that sits on top of the
 * CUDA library (include/svo_c.h) exposed to non-C++ callers (the Python parity tests, bench.py).
 *
 *   svo_vo_*     one stereo sequence driven through the reference-shaped C++ classes
 *                (include/stereo_vo/image_processor.hpp: ImageProcessor::process over
 *                FeatureTracker + BundleAdjuster bookkeeping) - the headless stand-in for the
 *                vo_node main loop, src/vo_node.cpp:112-119,141-144.
 *   svo_batch_*  many independent sequences per GPU (include/stereo_vo/batch_front_end.hpp).
 *   svo_camera_from_yaml   the camera YAML semantics of src/vo_node.cpp:82-98.
 *   svo_ba_*     the Ceres-free BundleAdjuster bookkeeping on its own (ids, window, world points).
 *   svo_kitti_*  the frame source of kitti_node (KITTI PNG pairs + pose file).
 *   svo_image_to_mono8, svo_stereo_queue_*   the image-message side of vo_node
 *                (cv_bridge::toCvCopy(..., MONO8), handle_images and its queue, the drain loop).
 *
 * Same conventions as svo_c.h: POD only, 0 on success / negative svo_status, nothing throws
 * across the boundary, *_last_error() returns the message of the last failure.
 */
#ifndef SVO_HOST_C_H_
#define SVO_HOST_C_H_

#include "svo_c.h"

#ifdef __cplusplus
extern "C" {
#endif

/* flat `key: value` YAML of config/ *.yaml -> the four numbers, narrowed to float */
SVO_API int svo_camera_from_yaml(const char* yaml_text, float* focal, float* cx, float* cy,
                                 float* baseline, float* frame_rate);

/* ---- single sequence through ImageProcessor::process --------------------------------- */
typedef struct svo_vo svo_vo;

/* Host PnP step (src/image_processor.cpp:75-92) supplied by the caller when the library is
 * built without OpenCV: n world points (xyz) and image points (xy), K row-major 3x3;
 * rvec/tvec: extrinsic guess in, solution out; rmat: row-major 3x3 rotation out; inliers:
 * room for n indices. Return 0 on success. */
typedef int (*svo_pnp_fn)(void* user, int32_t n, const float* world_xyz, const float* image_xy,
                          const float* K, double* rvec, double* tvec, double* rmat,
                          int32_t* inliers, int32_t* n_inliers);

enum { SVO_VO_NONE = 0, SVO_VO_TOO_FEW = 1, SVO_VO_FIRST = 2, SVO_VO_SKIP = 3, SVO_VO_KEYFRAME = 4 };

/* what process() did with the last pair; pointers stay valid until the next svo_vo_process() */
typedef struct svo_vo_record {
    int32_t decision;
    int32_t n_detected;      const svo_float2* detected;
    float av_parallax, percent_lost;
    int32_t n_tracked;       const svo_float2* tracked;       const uint64_t* tracked_ids;
    int32_t n_inliers;       const int32_t* inliers;
    float hmat[16];
    int32_t n_new_features;  const svo_float2* new_features;  const int16_t* disp16;
    int32_t n_new;           const int32_t* valid_idx;        const svo_float2* new_2d;
                             const svo_float3* new_3d;        const uint64_t* new_ids;
    int32_t n_tracker;       const svo_float2* tracker_pts;   const uint64_t* tracker_ids;
} svo_vo_record;

/* vo_node's construction sequence (src/vo_node.cpp:104-119): K from focal/cx/cy (float),
 * BundleAdjuster(window 5), FeatureTracker, ImageProcessor(K, ..., baseline, 30, 20) */
SVO_API int svo_vo_create(float focal, float cx, float cy, float baseline,
                          float min_feature_distance, float parallax_thresh, int32_t device,
                          svo_vo** out);
SVO_API void svo_vo_destroy(svo_vo* v);
SVO_API int svo_vo_set_pnp(svo_vo* v, svo_pnp_fn fn, void* user);
SVO_API int svo_vo_process(svo_vo* v, const uint8_t* left, const uint8_t* right, int32_t width,
                           int32_t height, int32_t pitch, double t);
SVO_API int svo_vo_last_record(svo_vo* v, svo_vo_record* out);
/* FeatureTracker::get_drawing (src/feature_tracker.cpp:85-91, published by vo_node.cpp:188): the
 * RGB overlay of the last keyframe's tracks, height rows of 3 * width bytes `out_pitch` apart.
 * SVO_ERR_STATE before the first frame. */
SVO_API int svo_vo_get_drawing(svo_vo* v, uint8_t* out_rgb, int32_t out_pitch, int32_t* width,
                               int32_t* height);
SVO_API const char* svo_vo_last_error(const svo_vo* v);

/* ---- many sequences per GPU ----------------------------------------------------------- */
typedef struct svo_batch svo_batch;

typedef struct svo_batch_summary {
    int64_t frames, keyframes, corners, tracked, new_points;
    int64_t right_images;    /* right images transferred with / for this step */
} svo_batch_summary;

SVO_API int svo_batch_create(const svo_params* p, int32_t device, int32_t n_seq, int32_t n_groups,
                             float focal, float cx, float cy, float baseline,
                             float parallax_thresh, svo_batch** out);
SVO_API void svo_batch_destroy(svo_batch* b);
SVO_API const char* svo_batch_last_error(const svo_batch* b);   /* b may be NULL (create) */
/* one frame of every sequence; see svo::BatchFrontEnd::step. left/right/next_*: n_seq pointers
 * (next_* may be NULL), poses: n_seq x 16 floats or NULL, summary may be NULL */
SVO_API int svo_batch_step(svo_batch* b, const uint8_t* const* left, const uint8_t* const* right,
                           int32_t pitch, int32_t on_device, const float* poses,
                           const uint8_t* const* next_left, const uint8_t* const* next_right,
                           svo_batch_summary* summary);
/* Pipelined form of svo_batch_step (svo::BatchFrontEnd::submit / wait): submit queues one frame
 * of every sequence and returns at once; wait blocks until the OLDEST submitted step is complete,
 * fills `summary` and makes that step the one svo_batch_get reads. At most 2 steps in flight
 * (SVO_ERR_STATE otherwise). Pointer arrays are copied at submit; the images must stay unchanged
 * until the matching wait returns. */
SVO_API int svo_batch_submit(svo_batch* b, const uint8_t* const* left, const uint8_t* const* right,
                             int32_t pitch, int32_t on_device, const float* poses,
                             const uint8_t* const* next_left, const uint8_t* const* next_right);
SVO_API int svo_batch_wait(svo_batch* b, svo_batch_summary* summary);

/* per-sequence results of the last step. Arrays: corners [max_corners], tracker_pts /
 * tracker_ids [max_features], new_2d / new_3d [max_corners]; any may be NULL */
SVO_API int svo_batch_get(svo_batch* b, int32_t seq, int32_t* decision, int32_t* n_corners,
                          svo_float2* corners, svo_track_stats* stats, int32_t* n_tracker,
                          svo_float2* tracker_pts, uint64_t* tracker_ids, int32_t* n_new,
                          svo_float2* new_2d, svo_float3* new_3d);
/* keyframe-stage taps of the last step for sequence seq (parity tests; zero counts unless the
 * decision was FIRST or KEYFRAME): detections that survived the new-vs-tracked dedupe
 * (src/image_processor.cpp:113-128) with their raw StereoBM values, and the indices of the
 * stereo-valid ones (:193-195). Arrays hold max_corners entries; any may be NULL */
SVO_API int svo_batch_get_keyframe(svo_batch* b, int32_t seq, int32_t* n_new_features,
                                   svo_float2* new_features, int16_t* disp16, int32_t* n_valid,
                                   int32_t* valid_idx);
/* Closed loop: the pose step of src/image_processor.cpp:75-92 runs on the GPU (svo_pnp_ransac)
 * instead of taking caller poses; call before the first step. iterations / reproj_error /
 * confidence: the arguments of cv::solvePnPRansac (reference: 100, 8, 0.99). */
SVO_API int svo_batch_set_closed_loop(svo_batch* b, int32_t on, int32_t iterations,
                                      float reproj_error, double confidence);
/* pose of sequence seq's last keyframe (row-major camera->world hmat, :130-134) and the inlier
 * count of the pose step of the last step (0 unless it was a closed-loop keyframe) */
SVO_API int svo_batch_get_pose(svo_batch* b, int32_t seq, float hmat[16], int32_t* n_inliers);
/* svo_set_blocking_sync for every engine group */
SVO_API int svo_batch_set_blocking_sync(svo_batch* b, int32_t on);
SVO_API int64_t svo_batch_kernel_launches(const svo_batch* b);
SVO_API int svo_batch_timer_begin(svo_batch* b);
SVO_API int svo_batch_timer_end(svo_batch* b, float* ms);
SVO_API int svo_batch_set_profiling(svo_batch* b, int32_t on);
SVO_API int svo_batch_get_profile(const svo_batch* b, double* ms, int64_t* counts);
/* engine of group g (device allocation helpers, debugging) */
SVO_API svo_engine* svo_batch_engine(svo_batch* b, int32_t group);

/* ---- KITTI odometry frame source (src/kitti_node.cpp:16-68 without ROS) ------------------ */
typedef struct svo_kitti svo_kitti;
/* data_path as in src/kitti_node.cpp:16 (trailing '/'), sequence "00".."21": parses
 * data_odometry_poses/dataset/poses/<seq>.txt and sizes the images from image_0/000000.png */
SVO_API int svo_kitti_open(const char* data_path, const char* sequence, svo_kitti** out);
SVO_API void svo_kitti_close(svo_kitti* k);
SVO_API const char* svo_kitti_last_error(const svo_kitti* k);   /* k may be NULL (open) */
SVO_API int svo_kitti_info(const svo_kitti* k, int64_t* n_frames, int32_t* width, int32_t* height);
/* frame i decoded into caller memory (8-bit grey, `pitch` bytes per row); pose12 receives the
 * 3x4 row-major [R | t] of the pose file (may be NULL) */
SVO_API int svo_kitti_read(svo_kitti* k, int64_t i, uint8_t* left, uint8_t* right, int32_t pitch,
                           double* pose12);
/* decode-ahead: frames first, first+1, ... decoded by a worker thread into `depth` >= 3 stereo
 * slots of `slab` (depth * 2 * pitch * height bytes, e.g. from svo_host_alloc). svo_kitti_next
 * returns 1 and the slot of the next frame (valid until two further calls), 0 at the end. */
SVO_API int svo_kitti_prefetch_start(svo_kitti* k, uint8_t* slab, int32_t depth, int32_t pitch,
                                     int64_t first);
/* the same with `threads` (1..64) decoder threads; frames still arrive in order. One thread
 * decodes ~90 KITTI-sized pairs per second (zlib inflate is 3/4 of it), 8 threads 690: a sequence
 * replayed faster than one thread decodes wants more (up to depth - 2 frames are decoded ahead). */
SVO_API int svo_kitti_prefetch_start_threads(svo_kitti* k, uint8_t* slab, int32_t depth,
                                             int32_t pitch, int64_t first, int32_t threads);
SVO_API int svo_kitti_next(svo_kitti* k, int64_t* index, const uint8_t** left, const uint8_t** right,
                           double* pose12);

/* ---- BundleAdjuster bookkeeping alone (src/bundle_adjuster.cpp:41-58,60-135,159-163) --------- */
/* The facade's Ceres-free BundleAdjuster (include/stereo_vo/bundle_adjuster.hpp) without an
 * ImageProcessor around it: id assignment, refcounts, sliding window and world-point store. Host
 * code; needs no GPU. tests/test_reference_ba.py replays, through these calls, the keyframe
 * sequences it also ran through the reference's own compiled bundle_adjuster.cpp. */
typedef struct svo_ba svo_ba;
SVO_API int svo_ba_create(int32_t window_size, svo_ba** out);
SVO_API void svo_ba_destroy(svo_ba* b);
SVO_API const char* svo_ba_last_error(const svo_ba* b);          /* b may be NULL (create) */
/* add_keyframe for a keyframe with n_tracked carried-over ids and n_new triangulated features:
 * new_ids (room for n_new) receives the id of every new feature that was kept, *n_kept their
 * number (n_new, or fewer when tracked + new exceeds the 400-feature limit: the reference
 * truncates there too but leaves zeros in front of the ids - SURVEY.md B6 - which this build
 * does not reproduce). SVO_ERR_INVALID for a tracked id that was never handed out. */
SVO_API int svo_ba_add_keyframe(svo_ba* b, const uint64_t* tracked_ids, int32_t n_tracked,
                                const svo_float3* new_3d, int32_t n_new, uint64_t* new_ids,
                                int32_t* n_kept);
SVO_API int svo_ba_get_world_points(svo_ba* b, const uint64_t* ids, int32_t n, svo_float3* out);
SVO_API int64_t svo_ba_num_features(const svo_ba* b);

/* ---- ROS image messages without ROS (src/vo_node.cpp:38-76,139-143; stereo_vo/image_msg.hpp) - */
/* the fields of sensor_msgs/Image that cv_bridge reads */
typedef struct svo_image_msg {
    double stamp;            /* header.stamp.toSec() */
    uint32_t height, width;
    const char* encoding;    /* sensor_msgs::image_encodings name: mono8, mono16, rgb8, bgr8, ... */
    uint8_t is_bigendian;
    uint32_t step;           /* bytes per row */
    const uint8_t* data;     /* step * height bytes */
    size_t size;
} svo_image_msg;
/* cv_bridge::toCvCopy(msg, "mono8") (src/vo_node.cpp:70-71) into caller memory: height rows of
 * width bytes, `pitch` apart. SVO_ERR_INVALID for a malformed message or an encoding cv_bridge
 * cannot turn into mono8 (8UC1, 32FC1, ...); the text is in svo_image_last_error(). */
SVO_API int svo_image_to_mono8(const svo_image_msg* msg, uint8_t* out, int32_t pitch);
SVO_API const char* svo_image_last_error(void);

/* handle_images + its queue<StereoPair> (src/vo_node.cpp:38-76): push converts a synchronised
 * pair to mono8 and queues it unless its left stamp is less than drop_time (the node: 0.05 s)
 * after the last queued pair's; *queued = 1 / 0 says which. */
typedef struct svo_stereo_queue svo_stereo_queue;
SVO_API int svo_stereo_queue_create(double drop_time, svo_stereo_queue** out);
SVO_API void svo_stereo_queue_destroy(svo_stereo_queue* q);
SVO_API const char* svo_stereo_queue_last_error(const svo_stereo_queue* q);   /* q may be NULL */
SVO_API int svo_stereo_queue_push(svo_stereo_queue* q, const svo_image_msg* left,
                                  const svo_image_msg* right, int32_t* queued);
SVO_API int32_t svo_stereo_queue_size(const svo_stereo_queue* q);
/* the node loop's drain (src/vo_node.cpp:140-143): every queued pair, oldest first, is handed to
 * fn and popped after fn returned 0; a non-zero return stops the drain with the pair still at
 * the front (SVO_ERR_STATE). *n_processed (nullable) counts the pairs popped. */
typedef int (*svo_stereo_pair_fn)(void* user, const uint8_t* left, const uint8_t* right,
                                  int32_t width, int32_t height, int32_t pitch_left,
                                  int32_t pitch_right, double t);
SVO_API int svo_stereo_queue_drain(svo_stereo_queue* q, svo_stereo_pair_fn fn, void* user,
                                   int32_t* n_processed);
/* the same with ImageProcessor::process of `v` as the consumer - vo_node's main loop body */
SVO_API int svo_vo_process_queue(svo_vo* v, svo_stereo_queue* q, int32_t* n_processed);

#ifdef __cplusplus
}
#endif
#endif /* SVO_HOST_C_H_ */
