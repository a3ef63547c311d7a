/* svo_c.h - C ABI of the B200-native stereo-VO front end (libsvo.so).
 *
 * Drop-in boundary for the pixel work inside the reference's ImageProcessor / FeatureTracker
 * (EpicDuckPotato/stereo_vo). The reference has no FFI layer of its own: its hot path is five
 * OpenCV calls inside two C++ classes. Each entry point below replaces one of those call sites
 * (file:line cited per function); include/stereo_vo/{image_processor,feature_tracker}.hpp keep
 * the reference's C++ signatures on top of this ABI, INTEGRATION.md shows the binding.
 *
 * Conventions
 *  - plain C: POD structs, raw pointers and sizes; no torch / OpenCV / Eigen types.
 *  - every function returns 0 on success or a negative svo_status; nothing throws across the
 *    ABI; svo_last_error() gives a human-readable string for the last failure of an engine.
 *  - caller pointers are never retained past a call: inputs are copied to the device
 *    (pinned staging + cudaMemcpyAsync), outputs are written to caller-owned arrays.
 *  - an engine owns one CUDA stream and the device state of `n_seq` independent sequences
 *    (camera streams). An engine is not thread-safe; distinct engines may be driven from
 *    different threads. Sequences of one engine are processed together by batched kernels.
 *  - images are 8-bit grey, row-major, `pitch` bytes per row, size fixed at svo_create().
 *  - there is no CPU fallback: if no CUDA device is usable every call fails with
 *    SVO_ERR_CUDA.
 *
 * Environment variables read by the libraries (once, at first use; none changes a result -
 * they select between kernels / schedules that the tests hold to the same output; they exist
 * for the tests and for tuning on other parts, and are NOT part of the stable ABI):
 *    SVO_CORNER_CPT = 1 | 2 | 4        columns per thread of corner_score_nms (default 2)
 *    SVO_CORNER_SEGMENTS = 1 .. 16     row segments per strip (default height / 62)
 *    SVO_CORNER_MIN_BLOCKS = 16 | 20   resident warps per SM the kernel is compiled for
 *    SVO_CORNER_FORCE_REPAIR = 1       every segment re-walks (exercises the repair path)
 *    SVO_STEREO_GENERIC = 1            sparse StereoBM: generic kernel even for (48, 21)
 *    SVO_SPLIT_STREAMS = 0 | 1         detection and tracking of a step on two streams
 *    SVO_LK_HANDOVER = 0               no backward -> forward hand-over in lk_track (svo_lk_handover)
 *    SVO_LK_REKEY = 1                  svo_tracker_set re-points the lk_track hand-over at the kept features (read by svo_create; off by default)
 *    libsvo_host.so (svo::BatchFrontEnd): SVO_BATCH_LAZY_RIGHT, SVO_BATCH_ZERO_COPY_RIGHT
 *    (how right images reach the device), SVO_BATCH_TRACE (per-step trace on stderr)
 */
#ifndef SVO_C_H_
#define SVO_C_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define SVO_API __attribute__((visibility("default")))
#else
#define SVO_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef enum svo_status {
    SVO_OK = 0,
    SVO_ERR_INVALID = -1,      /* bad argument */
    SVO_ERR_CUDA = -2,         /* CUDA runtime failure (see svo_last_error) */
    SVO_ERR_CAPACITY = -3,     /* a device buffer overflowed (candidates / features) */
    SVO_ERR_STATE = -4,        /* call sequence violated (e.g. track before init) */
    SVO_ERR_NOMEM = -5
} svo_status;

/* Algorithm constants. svo_default_params() fills the reference's hard-coded values. */
typedef struct svo_params {
    int32_t width, height;          /* image size of every frame given to this engine */
    /* cv::goodFeaturesToTrack(left, out, 300, 0.1, min_feature_distance)  image_processor.cpp:22 */
    int32_t max_corners;            /* 300 */
    double quality_level;           /* 0.1 */
    float min_distance;             /* 30 (vo_node.cpp:34) */
    int32_t max_candidates;         /* NMS survivor capacity per image; 0 = automatic */
    /* cv::StereoBM::create(16*3, 21)  image_processor.cpp:174 (+ OpenCV defaults) */
    int32_t num_disparities;        /* 48 */
    int32_t sad_window;             /* 21 */
    int32_t prefilter_cap;          /* 31 */
    int32_t texture_threshold;      /* 10 */
    int32_t uniqueness_ratio;       /* 15 */
    /* cv::calcOpticalFlowPyrLK(..., Size(21,21), 3, {COUNT+EPS,30,0.01}, 0, 1e-2)  feature_tracker.cpp:23-26 */
    int32_t lk_max_level;           /* 3 (window is fixed at 21x21) */
    int32_t lk_max_iters;           /* 30 */
    double lk_epsilon;              /* 0.01 */
    double lk_min_eig_threshold;    /* 1e-2 */
    /* FeatureTracker::track_features filter  feature_tracker.cpp:47,53 */
    float fb_max_error;             /* 2 px forward-backward */
    float max_parallax;             /* 200 px */
    int32_t max_features;           /* tracker capacity; 400 = bundle_adjuster.hpp:75 */
    /* debug: also store the cv::cornerMinEigenVal map (read by svo_debug_min_eigen_map); the
     * product path never materialises it */
    int32_t keep_eig_map;           /* 0 */
} svo_params;

typedef struct svo_engine svo_engine;   /* opaque */

typedef struct svo_float2 { float x, y; } svo_float2;
typedef struct svo_float3 { float x, y, z; } svo_float3;

typedef struct svo_track_stats {
    float av_parallax;      /* feature_tracker.cpp:63 */
    float percent_lost;     /* feature_tracker.cpp:64 */
    int32_t n_kept;
    int32_t n_attempted;
} svo_track_stats;

/* ---- lifetime ------------------------------------------------------------------------ */
SVO_API void svo_default_params(svo_params* p, int32_t width, int32_t height);
SVO_API int svo_create(const svo_params* p, int32_t device, int32_t n_seq, svo_engine** out);
SVO_API void svo_destroy(svo_engine* e);
SVO_API const char* svo_last_error(const svo_engine* e);   /* e may be NULL: last create() failure */
SVO_API int svo_get_params(const svo_engine* e, svo_params* out);
SVO_API int svo_num_levels(const svo_engine* e);            /* pyramid levels actually built */
/* kernels launched by this engine since creation (bench.py reports the delta per step) */
SVO_API int64_t svo_kernel_launches(const svo_engine* e);
/* Detections repeated on the spill lists so far. The candidate lists of the detector are sized
 * for ordinary images (a quarter of the pixels of a row segment; max_candidates survivors of
 * the quality threshold); an image with large plateaus of equal corner score can exceed them,
 * and cv::goodFeaturesToTrack has no such limit. svo_good_features_to_track, svo_step and
 * svo_step_device then repeat the detection of that sequence on lists with room for every pixel
 * (device memory allocated at the first overflow: 16 bytes per pixel, once per engine;
 * SVO_ERR_NOMEM if that fails) and return what OpenCV returns. svo_step_keyframe cannot - its
 * keyframe stage has consumed the corners by then - and reports SVO_ERR_CAPACITY. */
SVO_API int64_t svo_detector_spills(const svo_engine* e);
/* The hand-over inside svo_tracker_track / svo_step (feature_tracker.cpp:23-26,32-35 call
 * calcOpticalFlowPyrLK twice per frame): the forward pass of frame t+1 interpolates the same
 * image (frame t) around the same points (the tracked positions) as the backward pass of frame t
 * did, so the backward pass leaves its per-level window state (I, Ix, Iy as int16, the
 * spatial-gradient matrix, the min-eigenvalue verdict: 12 KB per feature) on the device and the
 * next forward pass takes it instead of recomputing it - only if the entry carries the serial of
 * the same pyramid build and the same point bit for bit, so results never depend on it.
 * After svo_tracker_set (a feature list chosen by the host) the look-ups of the next frame miss
 * and recompute; with SVO_LK_REKEY=1 in the environment of svo_create the call re-points the
 * look-up at the entries of the features it kept (a small kernel on a side stream), so they hit.
 * enable: 1 / 0 switch it on / off for this engine (on by default unless SVO_LK_HANDOVER=0 or
 * the buffers would exceed 8 GB), -1 leave as is. count: 1 / 0 count forward-pass look-ups and
 * hits from now on (a device atomic per feature), -1 leave as is. lookups / hits (nullable):
 * totals since counting was switched on (waits for the engine's stream). Returns SVO_OK, or
 * SVO_ERR_INVALID when enable = 1 is asked of an engine created without the buffers. */
SVO_API int svo_lk_handover(svo_engine* e, int enable, int count, int64_t* lookups, int64_t* hits);
/* device-side time of the last svo_step()/stage call's kernels, CUDA events on the engine stream */
SVO_API float svo_last_gpu_ms(const svo_engine* e);
/* ms[0] input copies, ms[1] kernels, ms[2] result copy of the last call (CUDA events) */
SVO_API int svo_last_timing(const svo_engine* e, float ms[3]);

/* How the calls of this engine wait for the GPU: 0 (default) spins on the stream (lowest
 * latency), 1 sleeps until the device raises an interrupt (cudaEventBlockingSync) - for hosts
 * with more driver threads than cores. No effect on results. */
SVO_API int svo_set_blocking_sync(svo_engine* e, int32_t on);

/* per-kernel profiling: when on, every kernel group of a call is bracketed by CUDA events on
 * the engine stream; totals (ms, launches of the group) accumulate per slot until reset by the
 * next svo_set_profiling() call. Slot names: svo_profile_name(0 .. svo_profile_slots()-1). */
SVO_API int svo_set_profiling(svo_engine* e, int32_t on);
SVO_API int svo_profile_slots(void);
SVO_API const char* svo_profile_name(int32_t slot);
SVO_API int svo_get_profile(const svo_engine* e, double* ms, int64_t* counts);
/* device-side stopwatch on the engine stream (CUDA events): begin synchronises first */
SVO_API int svo_timer_begin(svo_engine* e);
SVO_API int svo_timer_end(svo_engine* e, float* ms);

/* ---- stage calls: one reference call site each, single sequence, synchronous ---------- */

/* cv::goodFeaturesToTrack at src/image_processor.cpp:22. Uploads `img` as the current left
 * frame of `seq` (and builds its pyramid). out_xy holds max_corners entries. */
SVO_API int svo_good_features_to_track(svo_engine* e, int32_t seq, const uint8_t* img, int32_t pitch,
                               svo_float2* out_xy, int32_t* out_n);

/* ImageProcessor::triangulate_stereo, src/image_processor.cpp:165-208: StereoBM(48,21) at the
 * given features + Q reprojection with `pose` (4x4 row-major, camera->world). Outputs hold n
 * entries; valid_idx[k] is the input index of the k-th valid feature; disp16 is the raw 1/16 px
 * StereoBM value at EVERY input feature (n entries, -16 = rejected). */
SVO_API int svo_triangulate_stereo(svo_engine* e, int32_t seq, const uint8_t* left, const uint8_t* right,
                           int32_t pitch, const svo_float2* feats, int32_t n, const float pose[16],
                           float focal, float cx, float cy, float baseline,
                           svo_float3* out_3d, svo_float2* out_valid_2d, int32_t* out_valid_idx,
                           int16_t* out_disp16, int32_t* out_n_valid);

/* cv::StereoBM::create(16*3, 21)->compute(left, right, disp) at src/image_processor.cpp:173-176
 * as the FULL CV_16S map (the reference computes it and reads <= 300 pixels):
 * 1/16 px fixed point, -16 where StereoBM rejects or never evaluates a pixel. Uploads the pair
 * as the current frame of `seq`. out_disp16: height rows, out_pitch int16 elements apart
 * (>= width). Needs num_disparities 48 and sad_window 21, else SVO_ERR_INVALID. */
SVO_API int svo_stereo_bm_dense(svo_engine* e, int32_t seq, const uint8_t* left, const uint8_t* right,
                        int32_t pitch, int16_t* out_disp16, int32_t out_pitch);
/* Same maps for the stereo pairs already on the device (after svo_step / svo_set_right or any
 * stage call that uploaded both images), n sequences in one launch. out_disp16 may be NULL, or
 * hold NULL entries: those maps stay on the device (timing runs). */
SVO_API int svo_stereo_bm_dense_resident(svo_engine* e, int32_t n, const int32_t* seqs,
                                 int16_t* const* out_disp16, int32_t out_pitch);

/* FeatureTracker::init, src/feature_tracker.cpp:3-16 */
SVO_API int svo_tracker_init(svo_engine* e, int32_t seq, const uint8_t* img, int32_t pitch,
                     const svo_float2* feats, const uint64_t* ids, int32_t n);

/* FeatureTracker::track_features, src/feature_tracker.cpp:18-67 (LK forward, LK backward when
 * flow_back, filter, statistics; the tracker then holds the kept set and `img` as last_image) */
SVO_API int svo_tracker_track(svo_engine* e, int32_t seq, const uint8_t* img, int32_t pitch,
                      int32_t flow_back, float* av_parallax, float* percent_lost);

/* FeatureTracker::get_tracked_features, src/feature_tracker.cpp:69-72 (max_features entries) */
SVO_API int svo_tracker_get(svo_engine* e, int32_t seq, svo_float2* feats, uint64_t* ids, int32_t* n);

/* new-vs-tracked dedupe, src/image_processor.cpp:113-128: keep detection i iff no tracked point
 * within min_distance. out_* hold n_det entries. */
SVO_API int svo_dedupe(svo_engine* e, int32_t seq, const svo_float2* det, int32_t n_det,
               const svo_float2* tracked, int32_t n_tracked, float min_distance,
               svo_float2* out_xy, int32_t* out_keep_idx, int32_t* out_n);

/* taps of the last track call (parity tests): LK outputs before filtering, n_attempted entries.
 * For a point the FORWARD pass lost (status1 == 0) the backward pass is not run: reverse holds
 * the forward result and status2 is 0, where cv::calcOpticalFlowPyrLK would track it anyway
 * (src/feature_tracker.cpp:32-35). The reference never reads those entries (:45-47 tests status1
 * first), so the kept set is identical; only these two debug taps differ from cv2 there. */
SVO_API int svo_tracker_debug(svo_engine* e, int32_t seq, svo_float2* tracked, uint8_t* status1,
                      svo_float2* reverse, uint8_t* status2, int32_t* kept_idx, int32_t* n_attempted);

/* ---- component taps (parity tests / profiling) ----------------------------------------- */
/* eig map of cv::cornerMinEigenVal(img, 3, 3) for the current left frame of `seq`; out is
 * height x width floats, contiguous. SVO_ERR_STATE unless the engine keeps the map
 * (svo_params.keep_eig_map) */
SVO_API int svo_debug_min_eigen_map(svo_engine* e, int32_t seq, float* out);
/* pyramid level `level` of the current left frame; out is h_l x w_l bytes, contiguous */
SVO_API int svo_debug_pyramid_level(svo_engine* e, int32_t seq, int32_t level, uint8_t* out,
                            int32_t* w, int32_t* h);
/* sorted NMS candidates of the last detection (x, y), at most cap entries */
SVO_API int svo_debug_candidates(svo_engine* e, int32_t seq, int32_t* xy, int32_t cap, int32_t* n);

/* ---- fused, batched frame step (the throughput path used by ImageProcessor::process) ---- */

/* Phase A for n listed sequences: upload the stereo pair, build the left pyramid, detect
 * corners (image_processor.cpp:22) and - for sequences whose tracker is initialised - run
 * track_features (feature_tracker.cpp:18-67) against the previous frame. Tracking is skipped,
 * as in the reference, when fewer than 4 corners were found (image_processor.cpp:23-25).
 * Inputs are HOST pointers: left[i], right[i] for sequence seqs[i].
 * Outputs (host, caller-owned, may be NULL to skip a copy):
 *   corners   [n][max_corners]   n_corners [n]
 *   stats     [n]                (n_attempted == 0 when the tracker was not run)
 *   tracked   [n][max_features]  tracked_ids [n][max_features]   (kept set, n_kept entries) */
SVO_API int svo_step(svo_engine* e, int32_t n, const int32_t* seqs, const uint8_t* const* left,
             const uint8_t* const* right, int32_t pitch, svo_float2* corners, int32_t* n_corners,
             svo_track_stats* stats, svo_float2* tracked, uint64_t* tracked_ids);

/* Optional hint: start copying the stereo pairs of an UPCOMING svo_step() to the device now,
 * on a second stream, so the transfer overlaps the kernels of the step in flight (the
 * reference's vo_node keeps up to 5 queued StereoPairs, src/vo_node.cpp:36,73, so the next pair
 * is known ahead). Returns at once. A later svo_step() whose n, seqs, left, right and pitch are
 * identical consumes the staged frames instead of copying again; the host buffers must stay
 * unchanged until that step returns. At most two announcements may be pending (two staging
 * sets); a step for other frames than the oldest announcement copies normally, and an
 * announcement passed over twice is dropped. No effect on results. */
SVO_API int svo_prefetch(svo_engine* e, int32_t n, const int32_t* seqs, const uint8_t* const* left,
                 const uint8_t* const* right, int32_t pitch);

/* Same as svo_step but the frames are already resident in device memory (device pointers,
 * `pitch` bytes per row): the bench's kernel-only leg. */
SVO_API int svo_step_device(svo_engine* e, int32_t n, const int32_t* seqs, const uint8_t* const* d_left,
                    const uint8_t* const* d_right, int32_t pitch, svo_float2* corners,
                    int32_t* n_corners, svo_track_stats* stats, svo_float2* tracked,
                    uint64_t* tracked_ids);

/* Right image of the CURRENT frame for the listed sequences, for callers that gave svo_step()
 * no right images (right == NULL): the reference only ever reads the right image inside
 * triangulate_stereo, i.e. on keyframes (src/image_processor.cpp:34-39,137-142), so a caller
 * that knows the keyframe decision can skip ~3/4 of the right-image uploads. right[i] are host
 * pointers (device pointers when on_device), `pitch` bytes per row. Asynchronous: returns once
 * the copies are queued; the host images must stay unchanged until the next svo_keyframe() /
 * svo_triangulate_stereo() of the engine returns. */
/* 1 when kernels can read `p` directly (device memory, or pinned host memory mapped at the same
 * address, e.g. from svo_host_alloc): such images may be passed with on_device = 1, and are then
 * fetched by the ingest kernel itself - for pinned HOST memory that is a zero-copy read over
 * PCIe which does not queue behind bulk copies on the H2D copy engine. */
SVO_API int svo_device_can_read(const void* p);
SVO_API int svo_set_right(svo_engine* e, int32_t n, const int32_t* seqs,
                          const uint8_t* const* right, int32_t pitch, int32_t on_device);

/* Phase B for the sequences that became keyframes (image_processor.cpp:113-142): dedupe the
 * corners of the current frame against the PnP-inlier tracked points, StereoBM + reprojection of
 * the remaining new features with `poses[i]` (camera->world, 4x4 row major).
 *   inliers [n][max_features] with n_inliers[i] valid entries each (0 for the first keyframe)
 * Outputs per sequence i (stride max_corners):
 *   new_feats / n_new       detections that survived the dedupe, in order
 *   valid_2d, pts_3d, valid_idx (index into new_feats), n_valid; disp16 for every new feature */
SVO_API int svo_keyframe(svo_engine* e, int32_t n, const int32_t* seqs, const svo_float2* inliers,
                 const int32_t* n_inliers, const float* poses, float focal, float cx, float cy,
                 float baseline, svo_float2* new_feats, int32_t* n_new, svo_float2* valid_2d,
                 svo_float3* pts_3d, int32_t* valid_idx, int16_t* disp16, int32_t* n_valid);

/* svo_keyframe + svo_tracker_set in one call for drivers without a pose solver in the loop
 * (every feature that survived the last svo_step is an inlier): the inlier list is the
 * tracker's own device-resident feature set, and FeatureTracker::init(tracked + new features,
 * ids) (src/image_processor.cpp:144-162, ids as src/bundle_adjuster.cpp:98-116 assigns them)
 * runs on the device: the tracked features keep their ids, the stereo-valid new ones get
 * first_new_id[i], first_new_id[i] + 1, ... (clipped to max_features). Nothing but the
 * per-call descriptors travels to the GPU. Outputs as svo_keyframe. */
SVO_API int svo_keyframe_tracked(svo_engine* e, int32_t n, const int32_t* seqs, const float* poses,
                         const uint64_t* first_new_id, float focal, float cx, float cy,
                         float baseline, svo_float2* new_feats, int32_t* n_new,
                         svo_float2* valid_2d, svo_float3* pts_3d, int32_t* valid_idx,
                         int16_t* disp16, int32_t* n_valid);

/* What ImageProcessor::process does with a frame (src/image_processor.cpp:23-30,63-65). */
typedef enum svo_frame_decision {
    SVO_FRAME_TOO_FEW = 1,     /* fewer than 4 corners: frame ignored, tracker not advanced */
    SVO_FRAME_FIRST = 2,       /* no keyframe yet: this frame becomes the first one */
    SVO_FRAME_SKIP = 3,        /* tracked; parallax <= threshold and < 40 % of the tracks lost */
    SVO_FRAME_KEYFRAME = 4,    /* tracked; becomes a keyframe */
    SVO_FRAME_PENDING = 8      /* added to FIRST / KEYFRAME: right image was not on the device */
} svo_frame_decision;

/* svo_step and svo_keyframe_tracked in ONE call and one device round trip, for drivers without a
 * pose solver in the loop: after tracking, the decision above is taken on the device from the
 * statistics of this frame (parallax_thresh: src/vo_node.cpp:35), and the sequences that become
 * keyframes run the keyframe stage (dedupe against the tracked set, StereoBM at the new corners,
 * reprojection with poses[i], FeatureTracker::init with ids first_new_id[i] + k) at once.
 * on_device: left/right are device pointers (as svo_step_device). right[i] may be NULL; a
 * sequence that turns out to be a keyframe without a right image on the device comes back as
 * SVO_FRAME_FIRST/KEYFRAME + SVO_FRAME_PENDING with its tracker advanced but no keyframe stage
 * run: supply the image with svo_set_right and finish with svo_keyframe_tracked.
 * Step outputs as svo_step, keyframe outputs as svo_keyframe (valid where decision is FIRST or
 * KEYFRAME). */
SVO_API int svo_step_keyframe(svo_engine* e, int32_t n, const int32_t* seqs, const uint8_t* const* left,
                      const uint8_t* const* right, int32_t pitch, int32_t on_device,
                      const float* poses, const uint64_t* first_new_id, float parallax_thresh,
                      float focal, float cx, float cy, float baseline,
                      svo_float2* corners, int32_t* n_corners, svo_track_stats* stats,
                      svo_float2* tracked, uint64_t* tracked_ids, int32_t* decision,
                      svo_float2* new_feats, int32_t* n_new, svo_float2* valid_2d,
                      svo_float3* pts_3d, int32_t* valid_idx, int16_t* disp16, int32_t* n_valid);

/* ---- pose step ----------------------------------------------------- */
typedef struct svo_pnp_result {
    double rvec[3], tvec[3];        /* x_cam = R x_world + t, rvec = Rodrigues vector of R */
    double rmat[9];                 /* cv::Rodrigues(rvec), row major */
    int32_t n_inliers;
    int32_t ok;                     /* 0: fewer than 5 points, or no hypothesis with >= 5 inliers
                                       (cv::solvePnPRansac returns false): rvec/tvec = the guess */
    int32_t best_hypothesis;        /* index of the winning hypothesis, -1 if none */
    int32_t hypotheses_tried;       /* where OpenCV's confidence bound stopped the loop */
} svo_pnp_result;

/* cv::solvePnPRansac(world, image, K, no distortion, rvec, tvec, useExtrinsicGuess = true,
 * iterations, reproj_error, confidence, inliers) + cv::Rodrigues at src/image_processor.cpp:75-92,
 * for n independent point sets in one launch (one per sequence that became a keyframe).
 *   world [n][max_features] xyz, image [n][max_features] xy (pixels), counts[i] valid entries
 *   guess [n][6]: rvec, tvec going in (the reference keeps them from the previous keyframe)
 *   results [n]; inliers [n][max_features]: indices into the point set, ascending
 * iterations <= 128. Semantics kept from OpenCV: 5-point model, strictly-more-inliers rule,
 * confidence-bounded loop, inliers taken under the winning hypothesis, final pose = least
 * squares over the inliers. Samples and the minimal solver differ (csrc/pnp_ransac.cu), so the
 * pose agrees with cv2 to the tolerance stated in BASELINE.md rather than bit for bit. */
SVO_API int svo_pnp_ransac(svo_engine* e, int32_t n, const svo_float3* world,
                           const svo_float2* image, const int32_t* counts, float focal, float cx,
                           float cy, int32_t iterations, float reproj_error, double confidence,
                           const double* guess, svo_pnp_result* results, int32_t* inliers);

/* ---- bundle-adjustment consumer -------------------------------------- */
/* ReprojectionFactor::Evaluate (src/reprojection_factor.cpp:10-88) for the n_obs observations of a
 * window in one launch. poses [n_poses][7] = (qw, qx, qy, qz, tx, ty, tz), world wrt camera, as
 * BundleAdjuster::add_keyframe stores them (src/bundle_adjuster.cpp:63-70); points [n_points][3];
 * observation i = pixel obs[2i..2i+1] of point point_idx[i] seen from pose pose_idx[i] (the
 * residual blocks of src/bundle_adjuster.cpp:76-84,120-124). Outputs (any may be NULL):
 * residuals [n_obs][2], jac_pose [n_obs][2][7] and jac_point [n_obs][2][3], row major, in the
 * layout ceres hands to Evaluate. Focal length and principal point as in CameraInfo. */
SVO_API int svo_reprojection_evaluate(svo_engine* e, int32_t n_obs, const int32_t* pose_idx,
                                      const int32_t* point_idx, const double* obs,
                                      int32_t n_poses, const double* poses, int32_t n_points,
                                      const double* points, double focal, double cx, double cy,
                                      double* residuals, double* jac_pose, double* jac_point);

/* ---- debug overlay ---------------------------------------------------- */
/* FeatureTracker::draw_track, src/feature_tracker.cpp:74-83: image (8-bit grey, the tracker's
 * initial_image) -> out_rgb (height rows of 3 * width bytes, out_pitch bytes apart, R G B), with
 * a green cv::arrowedLine(from[i] -> to[i], CV_RGB(0, 255, 0), thickness 4) per feature. */
SVO_API int svo_draw_track(svo_engine* e, const uint8_t* image, int32_t pitch, const svo_float2* from,
                           const svo_float2* to, int32_t n, uint8_t* out_rgb, int32_t out_pitch);

/* FeatureTracker::init against the CURRENT frame of each listed sequence
 * (image_processor.cpp:51-52,148-162): feats/ids [n][max_features], counts[i] entries. */
SVO_API int svo_tracker_set(svo_engine* e, int32_t n, const int32_t* seqs, const svo_float2* feats,
                    const uint64_t* ids, const int32_t* counts);

/* ---- allocation-free dedupe ------------------------------------------------------------------ */
/* NO COUNTERPART IN THE REFERENCE: this entry point was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid allocation). It is svo_dedupe() - same rule, same fp32
 * arithmetic, same outputs in the same order - without an engine and without any allocation:
 * no cudaMalloc / cudaFree / cudaHostAlloc, no staging copy; one memset and one kernel are
 * queued on `stream` (a cudaStream_t, NULL = the default stream).
 *   device        CUDA device to run on (made current), or < 0 for the caller's current device
 *   det, tracked  n_det detections / n_tracked tracked points; out_xy (may be NULL),
 *                 out_keep_idx (may be NULL): room for n_det entries; out_n: one int32.
 *                 EVERY array is memory the device can dereference - device memory
 *                 (svo_device_alloc, cudaMalloc) or pinned host memory (svo_host_alloc); a
 *                 pageable host pointer is rejected with SVO_ERR_INVALID.
 *   wait          nonzero: synchronise `stream` before returning; zero: results are valid once
 *                 the caller has synchronised the stream (or the graph launch has completed)
 *   err, err_cap  optional buffer that receives a message when the return value is not SVO_OK
 * Limit: n_det <= SVO_DEDUPE_STATIC_MAX_DET; more is refused with SVO_ERR_CAPACITY (nothing is
 * queued). The scratch (one ticket word + one flag per detection, 65 792 bytes) is a static
 * device array inside libsvo.so: every GPU that loads the library holds it until the process
 * exits, and every call on a device shares it. The call clears what must start at zero itself;
 * calls that OVERLAP on one GPU (different streams without an event between them, parallel
 * graph branches) race on the scratch and give undefined output - order them.
 * The first call on a device looks the array's device address up (cudaGetSymbolAddress): make
 * that call outside stream capture. */
#define SVO_DEDUPE_STATIC_MAX_DET 16384
SVO_API int svo_dedupe_static(int32_t device, const svo_float2* det, int32_t n_det,
                              const svo_float2* tracked, int32_t n_tracked, float min_distance,
                              svo_float2* out_xy, int32_t* out_keep_idx, int32_t* out_n,
                              void* stream, int32_t wait, char* err, size_t err_cap);

/* device allocation helpers for callers that keep frames resident (bench kernel-only leg) */
SVO_API int svo_device_alloc(svo_engine* e, size_t bytes, void** out);
SVO_API int svo_device_free(svo_engine* e, void* p);
SVO_API int svo_device_upload(svo_engine* e, void* dst, const void* src, size_t bytes);
/* pinned host memory for frame buffers (asynchronous H2D from caller memory) */
SVO_API int svo_host_alloc(size_t bytes, void** out);
SVO_API int svo_host_free(void* p);

#ifdef __cplusplus
}
#endif
#endif /* SVO_C_H_ */
