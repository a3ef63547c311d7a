// Shared device/host definitions for the stereo-VO front-end kernels (sm_100a only).
//
// Data layout in HBM (all per engine, indexed by sequence `s`):
//   pyr    u8   [n_seq][2 slots][levels 0..L]   left-image pyramids, pitched rows (pitch % 128 == 0)
//   right  u8   [n_seq][H * pitch0]             right image of the current frame
//   eig    f32  [n_seq][H * eig_pitch]          Shi-Tomasi min-eigenvalue map (debug tap only)
//   strips u64  [n_seq][n_strips][n_seg][cap]   local maxima above the running threshold
//   cand   u64  [n_seq][max_candidates]         (ordered score bits << 32 | y*W+x), sorted
// Kernels never see host pointers; launchers take raw device pointers + a stream.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define SVO_MAX_LEVELS 6
#define SVO_LK_WIN 21            // reference: cv::Size(21, 21), src/feature_tracker.cpp:24

struct SvoLevel {
    int w, h, pitch;
    unsigned int off;            // byte offset of this level inside one pyramid slot
};

struct SvoPyrLayout {
    int n_levels;                // levels actually built (<= max_level + 1)
    SvoLevel lvl[SVO_MAX_LEVELS];
    unsigned int slot_bytes;     // bytes of one pyramid slot (all levels)
};

// per-call, per-sequence descriptor (uploaded from pinned memory before the kernels of a call)
struct SvoSeqCall {
    int seq;                     // sequence index inside the engine
    int cur_slot;                // pyramid slot that holds (or receives) the current frame
    int feat_buf;                // tracker feature buffer holding the input feature set
    int n_feat;                  // features in that buffer
    int n_initial;               // initial_features.size() of the reference tracker
    int n_pts;                   // generic point count of the call (stereo / dedupe)
    int n_aux;                   // generic secondary count (inliers for dedupe)
    int flags;
    float pose[16];              // camera -> world, row major
    // source of the current stereo pair for the ingest kernel (device pointers: either the
    // engine's contiguous H2D staging buffer or caller-owned device memory), rows src_pitch apart
    const uint8_t* src_left;
    const uint8_t* src_right;
    int src_pitch;
    int decision;                // svo_step_keyframe: per-frame decision taken on the device
    unsigned long long first_id; // svo_keyframe_tracked: id of the first new feature
};

// SvoSeqCall::flags
#define SVO_CALL_TRACKER 1       // the sequence's tracker is initialised
#define SVO_CALL_HAS_RIGHT 4     // a right image of the current frame is on the device
#define SVO_CALL_SKIP 8          // keyframe-stage kernels leave this sequence alone

__host__ __device__ __forceinline__ int svo_reflect101(int i, int n) {
    // BORDER_REFLECT_101 (gfedcb|abcdefgh|gfedcba) for |overshoot| < n
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}

// monotone float -> uint key (larger float <=> larger key), and back
__host__ __device__ __forceinline__ unsigned int svo_float_key(unsigned int bits) {
    return (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);
}
__host__ __device__ __forceinline__ unsigned int svo_key_float_bits(unsigned int key) {
    return (key & 0x80000000u) ? (key & 0x7fffffffu) : ~key;
}

#define SVO_CUDA_OK(expr)                                                   \
    do {                                                                    \
        cudaError_t _e = (expr);                                            \
        if (_e != cudaSuccess) return _e;                                   \
    } while (0)

// ------------------------------------------------------------------ launchers (one per .cu)
// Every launcher is asynchronous on `st`; `n_call` sequences are listed in `d_call`.

// img_serial (nullable): [n_seq][2] - the level-0 launch stamps `serial` on the slot it reads
cudaError_t svo_launch_pyr_down(uint8_t* pyr, const SvoPyrLayout& lay, size_t seq_stride,
                                const SvoSeqCall* d_call, int n_call, cudaStream_t st,
                                unsigned int* img_serial = nullptr, unsigned int serial = 0);

cudaError_t svo_launch_pyr_down_single(const uint8_t* src, int w, int h, int spitch,
                                       uint8_t* dst, int dpitch, cudaStream_t st);

// fused Shi-Tomasi score + 3x3 NMS: per-strip candidate lists (corner_score.cu)
int svo_corner_strips(int w);
int svo_corner_segments(int h);
int svo_corner_strip_cap(int h);          // per (strip, segment)
int svo_corner_strip_cap_full(int h);     // ... with room for every pixel (spill path)
size_t svo_corner_state_doubles();        // per (strip, segment)
struct SvoCornerCall {
    const uint8_t* pyr; size_t seq_stride;
    float* eig; int eig_pitch; size_t eig_stride;      // debug tap, nullptr: eig map not written
    unsigned int* eig_max; double quality;
    unsigned long long* strips; int n_strips, n_seg, strip_cap;
    int* strip_count; unsigned int* sub_max;
    unsigned int* seg_flag;                // [n_seq][n_strips][n_seg][2] epoch-tagged verdicts
    double* seg_state;
    unsigned int* ticket;                  // [n_seq], zero between launches
    unsigned int* work;                    // [2], zero between launches
    unsigned int epoch;                    // launch number on these buffers (> 0, < 2^30)
    const SvoSeqCall* d_call;
};
cudaError_t svo_launch_corner_score_nms(const SvoCornerCall& c, const SvoPyrLayout& lay,
                                        int n_call, cudaStream_t st);
cudaError_t svo_launch_filter_candidates(const unsigned long long* strips, int n_strips,
                                         int strip_cap, const int* strip_count,
                                         const unsigned int* eig_max, double quality,
                                         unsigned long long* cand, int cand_stride, int max_cand,
                                         int* cand_count, const SvoSeqCall* d_call, int n_call,
                                         cudaStream_t st);

int svo_sort_capacity_pad(int max_cand);
int svo_sort_launch_count(int max_cand);
cudaError_t svo_sort_init();
cudaError_t svo_launch_sort_candidates(unsigned long long* cand, int max_cand, int* cand_count,
                                       const unsigned long long* strips, int n_strips,
                                       int strip_cap, const int* strip_count,
                                       const unsigned int* eig_max, double quality,
                                       const SvoSeqCall* d_call, int n_call, cudaStream_t st);

int svo_greedy_grid_ints(int w, int h, float min_distance);
cudaError_t svo_greedy_init();
cudaError_t svo_launch_greedy_select(const unsigned long long* cand, int max_cand,
                                     const int* cand_count, int w, int h, float min_distance,
                                     int max_corners, float2* corners, int* n_corners,
                                     int* grid_scratch, size_t grid_stride,
                                     const SvoSeqCall* d_call, int n_call, cudaStream_t st);

struct SvoStereoParams {
    int ndisp, win, cap, texture, uniq;
};

cudaError_t svo_launch_stereo_bm_sparse(const uint8_t* pyr, const SvoPyrLayout& lay,
                                        size_t seq_stride, const uint8_t* right,
                                        size_t right_stride, const float2* pts, int pts_stride,
                                        short* disp16, SvoStereoParams sp, int max_pts,
                                        const SvoSeqCall* d_call, int n_call, cudaStream_t st);

bool svo_stereo_dense_supported(const SvoStereoParams& sp);
cudaError_t svo_launch_stereo_bm_dense(const uint8_t* pyr, const SvoPyrLayout& lay,
                                       size_t seq_stride, const uint8_t* right,
                                       size_t right_stride, short* disp, int dpitch,
                                       size_t disp_stride, SvoStereoParams sp,
                                       const SvoSeqCall* d_call, int n_call, cudaStream_t st);

cudaError_t svo_launch_reproject(const float2* pts, const short* disp16, int pts_stride,
                                 float focal, float cx, float cy, float baseline, float2* valid2d,
                                 float* pts3d, int* valid_idx, int* n_valid,
                                 const SvoSeqCall* d_call, int n_call, cudaStream_t st);

cudaError_t svo_launch_dedupe(const float2* det, const int* n_det, int det_stride,
                              const float2* inl, int inl_stride, int inl_buffered,
                              float min_distance, float2* out,
                              int* keep_idx, int* n_out, int out_stride,
                              SvoSeqCall* d_call, unsigned* ticket, int n_call, cudaStream_t st);

// result block of pnp_ransac (layout of svo_pnp_result in include/svo_c.h)
struct SvoPnpResult {
    double rvec[3], tvec[3], rmat[9];
    int n_inliers, ok, best_hypothesis, hypotheses_tried;
};
cudaError_t svo_launch_pnp_ransac(const float* world, const float2* image, int stride,
                                  const int* counts, const double* guess, float focal, float cx,
                                  float cy, int iterations, float reproj_error, double confidence,
                                  unsigned salt, SvoPnpResult* results, int* inliers, int n_call,
                                  cudaStream_t st);

cudaError_t svo_launch_reprojection_factor(const double* poses, const double* points,
                                           const int* pose_idx, const int* point_idx,
                                           const double* obs, int n_obs, double f, double cx,
                                           double cy, double* resid, double* jac_pose,
                                           double* jac_point, cudaStream_t st);

cudaError_t svo_launch_draw_track(const uint8_t* gray, int gpitch, const float2* from,
                                  const float2* to, int n, uint8_t* rgb, int rpitch, int w, int h,
                                  int thickness, cudaStream_t st);

struct SvoLkParams {
    int max_iters;
    double eps2;                 // criteria.epsilon squared (OpenCV squares it in double)
    double min_eig;              // minEigThreshold (double in the OpenCV signature)
    int grid_pts;                // blocks per sequence to launch (0 = max_pts)
};

// Per-point window state handed from the backward pass of one frame to the forward pass of the
// next (lk_track.cu). All pointers null: no hand-over.
struct SvoLkCacheHdr {
    unsigned int serial;         // pyramid build the I patches came from (0: never written)
    float ptx, pty;              // the key-point, bit for bit
    unsigned int status;         // one byte per level: 0 not reached, 1 stored, 3 min-eigenvalue reject
    float4 a[4];                 // A11, A12, A22 (scaled), 1 / D per level (lk_track: <= 4 levels)
};
struct SvoLkCache {
    uint4* data;                 // [n_seq * max_pts][levels][3][64 threads]: packed int16 I, Ix, Iy
    SvoLkCacheHdr* hdr;          // [n_seq * max_pts]
    const unsigned int* img_serial;   // [n_seq][2 slots]: stamped by every pyramid build
    const int* kept_idx;         // [n_seq][max_pts]: tracked-array index of tracker feature k
    unsigned long long* stats;   // nullable: [0] forward-pass look-ups, [1] hits
};

// svo_tracker_set: sequences whose feature list the host just replaced (lk_rekey_kernel)
struct SvoLkRekey {
    int n;                       // entries used (<= 32)
    int seq[32], slot[32];       // sequence, pyramid slot of its current image
    int feat_buf[32], count[32]; // tracker feature buffer holding the new list, its length
};
cudaError_t svo_launch_lk_rekey(const float2* feat, int max_pts,
                                const SvoLkRekey& rk, const SvoLkCache& cache, int* kept_idx,
                                cudaStream_t st);

// direction 0: prev slot -> cur slot ; direction 1: cur slot -> prev slot
cudaError_t svo_launch_lk_track(const uint8_t* pyr, const SvoPyrLayout& lay, size_t seq_stride,
                                const float2* in_pts, size_t in_stride, int in_buffered,
                                float2* out_pts, unsigned char* status, int out_stride,
                                int direction, SvoLkParams lp, int max_pts, const int* gate,
                                int gate_min, const unsigned char* skip_status,
                                const SvoSeqCall* d_call, int n_call, cudaStream_t st,
                                SvoLkCache cache = SvoLkCache{nullptr, nullptr, nullptr, nullptr, nullptr});

struct SvoTrackStats {
    float av_parallax;
    float percent_lost;
    int n_kept;
    int n_attempted;
};

cudaError_t svo_launch_track_filter(float2* feat, unsigned long long* ids,
                                    float2* initial, size_t feat_stride,
                                    const float2* tracked, const unsigned char* status1,
                                    const float2* reverse, const unsigned char* status2,
                                    int out_stride, int flow_back, float fb_max, float max_parallax,
                                    int* kept_idx, float* kept_parallax, SvoTrackStats* stats,
                                    int max_features, const int* gate, int gate_min,
                                    const SvoSeqCall* d_call, int n_call, cudaStream_t st);
