// batch_front_end.hpp - throughput driver for many independent stereo sequences on one GPU
// (BASELINE configs[3]: 64 KITTI-shaped sequences sharded over 1/2/4/8 B200).
//
// The reference processes one camera stream per vo_node process, single-threaded
// (src/vo_node.cpp:139-227). Here S sequences of one GPU are split into G groups; every group
// owns one svo_engine (one CUDA stream, batched kernels over its sequences) and one host worker
// thread that runs the per-frame control flow of ImageProcessor::process for its sequences:
//   svo_step -> keyframe decision (src/image_processor.cpp:23-25,63-65) -> svo_keyframe ->
//   id assignment (src/bundle_adjuster.cpp:98-116) -> svo_tracker_set.
// Groups run concurrently, so the host work and the launch-latency-bound small kernels of one
// group overlap the heavy kernels of the others, and (host frames) the H2D copies of the next
// step overlap the kernels of the current one (svo_prefetch). Sequences never interact: there
// is no collective, on one GPU or across GPUs.
//
// Two ways to get a keyframe's pose (src/image_processor.cpp:75-92):
//   * teacher-forced (default): a keyframe keeps every tracked feature as inlier and takes the
//     camera->world pose the caller supplies for that frame;
//   * closed loop (set_closed_loop): the pose step runs on the GPU (svo_pnp_ransac, batched over
//     the group's keyframes) on the tracked features and the world points this driver stored
//     when it handed out their ids (BundleAdjuster::get_world_points without Ceres: un-refined
//     points, src/bundle_adjuster.cpp:159-163); inliers, hmat = [R^T | -R^T t], dedupe, stereo
//     and tracker re-init then follow the reference line by line (:95-162).
#pragma once
#include <condition_variable>
#include <cstdint>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "svo_c.h"

namespace svo {

struct BatchOptions {
    svo_params params;            // svo_default_params() + overrides; width/height required
    int device = 0;
    int n_seq = 1;
    int n_groups = 1;
    float focal = 0.f, cx = 0.f, cy = 0.f, baseline = 0.f;
    float parallax_thresh = 20.f; // src/vo_node.cpp:35
};

enum BatchDecision : int32_t { BATCH_TOO_FEW = 1, BATCH_FIRST = 2, BATCH_SKIP = 3, BATCH_KEYFRAME = 4 };

struct BatchSummary {             // totals over all sequences of one step
    int64_t frames = 0, keyframes = 0, corners = 0, tracked = 0, new_points = 0;
    int64_t right_images = 0;     // right images transferred (all, or predicted + on demand)
};

class BatchFrontEnd {
  public:
    explicit BatchFrontEnd(const BatchOptions& opt);
    ~BatchFrontEnd();
    BatchFrontEnd(const BatchFrontEnd&) = delete;
    BatchFrontEnd& operator=(const BatchFrontEnd&) = delete;

    // One frame of every sequence. left/right: n_seq image pointers (host memory, or device
    // memory when on_device), `pitch` bytes per row. poses: n_seq x 16 floats (row-major
    // camera->world, used only by sequences that become keyframes; a sequence's FIRST keyframe
    // always takes the identity, as src/image_processor.cpp:34-39 does) or nullptr for identity.
    // next_left/next_right: optional host frames of the following step, copied ahead while
    // this step computes. Throws std::runtime_error on a C-ABI failure.
    BatchSummary step(const uint8_t* const* left, const uint8_t* const* right, int pitch,
                      bool on_device, const float* poses, const uint8_t* const* next_left = nullptr,
                      const uint8_t* const* next_right = nullptr);

    // Pipelined form: submit() queues one frame of every sequence and returns at once, wait()
    // blocks until the OLDEST submitted step is complete and makes its results current. Up to
    // kPipeDepth steps may be in flight. Groups then run free of each other: while one group is
    // in the latency-bound tail of step k (keyframe stage, tracker re-init, host bookkeeping),
    // others already run the heavy kernels of step k+1. The pointer arrays are copied at
    // submit(); the IMAGES must stay unchanged until the matching wait() returns (and, for
    // next_left/next_right, until the step that consumes them has been waited for).
    static constexpr int kPipeDepth = 2;
    void submit(const uint8_t* const* left, const uint8_t* const* right, int pitch, bool on_device,
                const float* poses, const uint8_t* const* next_left = nullptr,
                const uint8_t* const* next_right = nullptr);
    BatchSummary wait();
    int in_flight() const { return (int)(n_submitted_ - n_waited_); }

    int n_seq() const { return opt_.n_seq; }
    int n_groups() const { return (int)groups_.size(); }
    int max_corners() const { return opt_.params.max_corners; }
    int max_features() const { return opt_.params.max_features; }

    // results of the last completed (step() / wait()) step for sequence s
    int32_t decision(int s) const { return res(s).decision; }
    int32_t n_corners(int s) const { return res(s).n_corners; }
    const svo_float2* corners(int s) const;
    const svo_track_stats& stats(int s) const { return res(s).stats; }
    // tracker set after the step (what FeatureTracker holds): points, ids, count
    int32_t tracker_count(int s) const { return res(s).n_tracker; }
    const svo_float2* tracker_points(int s) const;
    const uint64_t* tracker_ids(int s) const;
    // keyframe outputs of the step (valid when decision is FIRST or KEYFRAME)
    int32_t n_new(int s) const { return res(s).n_new; }
    const svo_float2* new_2d(int s) const;
    const svo_float3* new_3d(int s) const;
    // taps of the keyframe stage (parity tests): detections that survived the dedupe
    // (src/image_processor.cpp:113-128), their raw StereoBM values (1/16 px, -16 = rejected) and
    // the indices of the stereo-valid ones (:193-195; n_valid may exceed n_new when the
    // 400-feature clip of src/bundle_adjuster.cpp:75 applies)
    int32_t n_new_features(int s) const { return res(s).n_new_features; }
    int32_t n_valid(int s) const { return res(s).n_valid; }
    const svo_float2* new_features(int s) const;
    const int16_t* disp16(int s) const;
    const int32_t* valid_idx(int s) const;

    // Closed-loop pose step: call before the first step. Arguments of cv::solvePnPRansac at
    // src/image_processor.cpp:80.
    void set_closed_loop(bool on, int iterations = 100, float reproj_error = 8.f,
                         double confidence = 0.99);
    bool closed_loop() const { return closed_loop_; }
    // pose of the sequence's last keyframe (row-major camera->world, src/image_processor.cpp:
    // 130-134) and the inlier count of its pose step (closed loop only)
    const float* pose(int s) const { return seq_[s].hmat; }
    int32_t n_inliers(int s) const { return res(s).n_inliers; }

    // waits of every engine group sleep instead of spinning (svo_set_blocking_sync)
    void set_blocking_sync(bool on);

    int64_t kernel_launches() const;
    // device-side stopwatch over all groups: begin() synchronises every engine stream and
    // starts a CUDA-event timer on each; end() returns the longest elapsed time (ms)
    void timer_begin();
    float timer_end();
    void set_profiling(bool on);
    // sums over groups, indexed by svo_profile_name() slot
    void get_profile(std::vector<double>& ms, std::vector<int64_t>& counts) const;
    svo_engine* engine(int group) const { return groups_[group]->e; }

  private:
    struct SeqResult {                     // per pipeline slot
        int32_t decision = 0, n_corners = 0, n_tracker = 0, n_new = 0;
        int32_t n_new_features = 0, n_valid = 0, n_inliers = 0;
        svo_track_stats stats{};
    };
    struct SeqState {
        int group = 0, local = 0;          // engine and index inside it
        bool have_kf = false;
        uint64_t next_id = 0;
        int32_t n_tracker = 0;             // tracker set size carried from step to step
        float par_last = 0.f, par_rate = 8.f;   // keyframe predictor: last av_parallax, growth/step
        // closed loop: world point of every id handed out (ids are a running counter from 0),
        // the extrinsic guess carried from keyframe to keyframe (rvec, tvec), last keyframe pose
        std::vector<svo_float3> world;
        double rt[6] = {0, 0, 0, 0, 0, 0};
        float hmat[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
        SeqResult res[kPipeDepth];
    };
    struct Group {
        svo_engine* e = nullptr;
        std::vector<int> seqs;             // global sequence indices, local order
        std::thread worker;
        uint64_t done = 0;                 // steps this group has completed
        // scratch (all sized for the group's sequences)
        std::vector<int32_t> local_ids, kf_local, n_inl, n_newv, n_valid, counts, n_corn;
        std::vector<const uint8_t*> lp, rp, nlp, nrp, srp;
        std::vector<int32_t> miss_local;
        std::vector<char> pred_cur, pred_next;   // right image travels with this / the next step
        bool have_pred = false;
        int64_t pred_sent = 0, pred_hits = 0, pred_miss = 0;
        std::vector<svo_track_stats> stats;
        std::vector<svo_float2> corners, tracked, inl, new_feats, valid_2d, feats;
        std::vector<svo_float3> pts_3d;
        std::vector<uint64_t> tracked_ids, ids, first_id;
        std::vector<int32_t> valid_idx;
        std::vector<int16_t> disp16;
        std::vector<float> poses;
        // closed loop: pose-step inputs / outputs of the group's keyframes
        std::vector<svo_float3> pnp_world;
        std::vector<svo_float2> pnp_image;
        std::vector<int32_t> pnp_counts, pnp_inl, pnp_kf;
        std::vector<double> pnp_guess;
        std::vector<svo_pnp_result> pnp_res;
        double phase_ms[12] = {0};         // SVO_BATCH_TRACE=1: accumulated per-phase times
        int64_t phase_calls = 0;
    };
    struct Job {                           // one submitted step (pointer arrays owned)
        std::vector<const uint8_t*> left, right, next_left, next_right;
        std::vector<float> poses;
        bool has_right = false, has_next = false, has_next_right = false, has_poses = false;
        int pitch = 0;
        bool on_device = false;
        std::vector<BatchSummary> sum;     // per group
        std::vector<std::string> error;    // per group
        int pending = 0;                   // groups still working on it
    };
    void run_group(Group& g, Job& job, int slot, int gi);
    void run_keyframes_closed_loop(Group& g, int slot, int nk, BatchSummary& sum);
    void worker_loop(Group* g, int gi);
    const SeqResult& res(int s) const { return seq_[s].res[cur_slot_]; }

    BatchOptions opt_;
    bool trace_ = false;           // SVO_BATCH_TRACE=1: per-phase host/device times on stderr
    bool zero_copy_right_ = true;  // keyframe right images read in place from mapped pinned memory
    bool lazy_right_ = true;       // upload right images for keyframes only (SVO_BATCH_LAZY_RIGHT=0: always)
    bool closed_loop_ = false;
    int pnp_iterations_ = 100;
    float pnp_reproj_ = 8.f;
    double pnp_confidence_ = 0.99;
    std::vector<std::unique_ptr<Group>> groups_;
    std::vector<SeqState> seq_;
    // per-sequence result storage (global index), one set per pipeline slot
    std::vector<svo_float2> corners_[kPipeDepth], tracker_pts_[kPipeDepth], new_2d_[kPipeDepth];
    std::vector<svo_float2> new_feats_[kPipeDepth];
    std::vector<int16_t> disp16_[kPipeDepth];
    std::vector<int32_t> valid_idx_[kPipeDepth];
    std::vector<svo_float3> new_3d_[kPipeDepth];
    std::vector<uint64_t> tracker_ids_[kPipeDepth];
    Job jobs_[kPipeDepth];
    std::mutex mu_;
    std::condition_variable cv_start_, cv_done_;
    uint64_t n_submitted_ = 0, n_waited_ = 0;
    int cur_slot_ = 0;                     // slot of the last step handed back by wait()
    bool stop_ = false;
};

}  // namespace svo
