// libsvo.so: the C ABI of include/svo_c.h on top of the sm_100a kernels in this directory.
// One engine = one CUDA stream + the device state of n_seq independent stereo sequences.
// Host-side responsibilities only: buffer layout, per-call descriptors, H2D/D2H staging,
// tracker slot bookkeeping (which pyramid slot is FeatureTracker::last_image), error strings.
// There is no CPU implementation of any stage in here: without a CUDA device every call fails.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/svo_c.h"
#include "svo_common.cuh"

#define SVO_PROF_SLOTS 14
static const char* const kProfNames[SVO_PROF_SLOTS] = {
    "pyr_down_u8", "corner_score_nms", "ingest", "sort_candidates",
    "greedy_min_distance", "lk_track_fwd", "lk_track_bwd", "track_filter_compact",
    "pack_results", "dedupe_new_features", "stereo_bm_sparse", "reproject_q", "stereo_bm_dense",
    "pnp_ransac"};
enum { P_PYR, P_SCORE, P_INGEST, P_SORT, P_GREEDY, P_LKF, P_LKB, P_FILTER, P_PACK, P_DEDUPE,
       P_STEREO, P_REPROJ, P_DENSE, P_PNP };

namespace {

constexpr int CALL_RING = 8;

struct SeqState {
    int last_slot = -1;       // pyramid slot that is FeatureTracker::last_image (-1: none)
    int cur_slot = 0;         // slot of the most recently uploaded left frame
    bool have_frame = false;
    bool have_right = false;
    bool tracker_init = false;
    int feat_buf = 0;
    int n_feat = 0;
    int n_initial = 0;
    int dbg_attempted = 0;    // n_attempted of the last track call
};

// header of the packed per-sequence result block of a step
struct StepHeader {
    int n_corners;
    int cand_count;
    SvoTrackStats stats;
};
struct KeyHeader {
    int n_new;
    int n_valid;
    int decision;                // svo_step_keyframe: svo_frame_decision taken on the device
    int n_tracked;               // ... and the tracked features the new tracker set starts with
};

template <typename T>
cudaError_t dalloc(T** p, size_t count) {
    return cudaMalloc(reinterpret_cast<void**>(p), count * sizeof(T) + 256);
}

}  // namespace

struct svo_engine {
    svo_params p;
    int device = 0;
    int n_seq = 0;
    cudaStream_t st = nullptr;
    std::string err;
    SvoPyrLayout lay;
    size_t seq_stride = 0;
    size_t right_stride = 0;
    int pitch0 = 0;
    int eig_pitch = 0;
    size_t eig_stride = 0;
    int max_cand = 0;
    size_t grid_stride = 0;

    uint8_t* d_pyr = nullptr;
    uint8_t* d_right = nullptr;
    short* d_dense = nullptr;         // [n_seq][height][width] CV_16S maps of svo_stereo_bm_dense (allocated on first use)
    uint8_t* d_stage = nullptr;       // [2 sets][n_seq][2] contiguous (unpitched) H2D landing buffers
    size_t stage_img = 0;             // bytes of one staged image (width*height rounded to 256)
    size_t stage_set = 0;             // bytes of one staging set
    // svo_prefetch(): stereo pairs of upcoming steps copied ahead on a second stream while the
    // kernels of the current step run (FIFO, one entry per staging set)
    struct Prefetch {
        bool valid = false;
        int set = 0, n = 0, pitch = 0;
        std::vector<int32_t> seqs;
        std::vector<const uint8_t*> left, right;
        cudaEvent_t done = nullptr;
        int skipped = 0;              // steps that passed this entry over (other frames)
    };
    Prefetch pf[2];
    int pf_head = 0, pf_count = 0;    // oldest pending entry, number pending
    int cur_set = 0;                  // staging set of the uploads of the current call
    cudaStream_t st_copy = nullptr;
    // svo_step runs detection (corner score -> repair -> sort -> greedy) on a second stream next
    // to pyramid + LK tracking: the two halves of a frame only meet at the tracker gate
    // (image_processor.cpp:23-25) and in the result block
    cudaStream_t st_det = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    bool split_streams = true;        // SVO_SPLIT_STREAMS=0: one stream
    bool blocking_sync = false;       // svo_set_blocking_sync
    cudaEvent_t ev_block = nullptr;   // cudaEventBlockingSync: waits sleep instead of spinning
    float* d_eig = nullptr;
    unsigned int* d_eig_max = nullptr;
    unsigned long long* d_cand = nullptr;
    int* d_cand_count = nullptr;
    unsigned long long* d_strips = nullptr;   // [n_seq][n_strips][n_seg][strip_cap]
    int* d_strip_count = nullptr;             // [n_seq][n_strips * n_seg]
    int n_strips = 0, n_seg = 0, strip_cap = 0;
    // spill path of the detector (plateau images): lists with room for EVERY pixel of one
    // sequence, allocated at the first overflow; spill_seq = the sequence whose sorted
    // candidates they hold
    unsigned long long* d_strips_big = nullptr;   // [n_strips][n_seg][strip_cap_big]
    unsigned long long* d_cand_big = nullptr;     // [svo_sort_capacity_pad(max_cand_big)]
    int strip_cap_big = 0, max_cand_big = 0, spill_seq = -1;
    int64_t spills = 0;
    // corner_score_nms: per (strip, segment) maximum, wrong-start flag, true sums of a flagged start
    unsigned int* d_sub_max = nullptr;
    unsigned int* d_seg_flag = nullptr;       // [..][2] epoch-tagged start verdicts
    double* d_seg_state = nullptr;
    unsigned int* d_cs_ticket = nullptr;
    unsigned int* d_cs_work = nullptr;        // [2] work-item dispenser of corner_score_nms
    unsigned int cs_epoch = 0;                // corner_score_nms launches on these buffers
    float2* d_corners = nullptr;
    int* d_ncorners = nullptr;
    int* d_grid = nullptr;
    float2* d_feat = nullptr;
    unsigned long long* d_ids = nullptr;
    float2* d_initial = nullptr;
    float2* d_tracked = nullptr;
    unsigned char* d_status1 = nullptr;
    float2* d_reverse = nullptr;
    unsigned char* d_status2 = nullptr;
    int* d_kept_idx = nullptr;
    // lk_track hand-over (backward pass of frame t -> forward pass of frame t + 1)
    uint4* d_lk_data = nullptr;
    SvoLkCacheHdr* d_lk_hdr = nullptr;
    unsigned int* d_img_serial = nullptr;
    unsigned int pyr_serial = 0;
    bool lk_handover = false;            // buffers present and switched on
    // svo_tracker_set re-points the hand-over at the features the host kept (lk_rekey_kernel),
    // on a high-priority stream of its own: a launch on the engine stream would queue behind
    // the other engine groups' grids on every keyframe step
    bool lk_rekey = false;
    bool rekey_pending = false;          // ev_rekey not yet waited for by the engine stream
    cudaStream_t st_rekey = nullptr;
    cudaEvent_t ev_feat = nullptr, ev_rekey = nullptr;
    bool lk_count = false;
    unsigned long long* d_lk_stats = nullptr;
    float* d_kept_par = nullptr;
    SvoTrackStats* d_stats = nullptr;
    float2* d_inl = nullptr;
    float2* d_pts = nullptr;      // stereo input points / deduped new features
    int* d_keep_idx = nullptr;
    int* d_nnew = nullptr;
    unsigned* d_ticket = nullptr;     // dedupe: blocks of a sequence that have finished (0 between calls)
    short* d_disp16 = nullptr;
    float2* d_valid2d = nullptr;
    float* d_pts3d = nullptr;
    int* d_valid_idx = nullptr;
    int* d_nvalid = nullptr;
    // svo_pnp_ransac (allocated on first use): point sets, guesses, results
    float* d_pnp_world = nullptr;
    float2* d_pnp_img = nullptr;
    int* d_pnp_counts = nullptr;
    double* d_pnp_guess = nullptr;
    SvoPnpResult* d_pnp_res = nullptr;
    int* d_pnp_inl = nullptr;
    unsigned pnp_calls = 0;
    uint8_t* d_draw = nullptr;        // svo_draw_track: RGB image, grey image, arrow end points
    unsigned char* d_ba = nullptr;    // svo_reprojection_evaluate scratch (grown on demand)
    size_t ba_bytes = 0;
    unsigned char* d_out = nullptr;   // packed result blocks
    size_t step_block = 0, key_block = 0, out_bytes = 0;
    int pts_cap = 0;                  // per-sequence capacity of d_pts / stereo outputs

    SvoSeqCall* h_call = nullptr;     // pinned + device-mapped ring [CALL_RING][n_seq]
    // small host->device inputs (descriptors, inlier / feature lists) are PULLED by a kernel
    // straight out of mapped pinned memory instead of going through cudaMemcpyAsync: the H2D
    // copy engine is a FIFO shared by all streams, and a 30 MB svo_prefetch() queued on it
    // would hold a 7 KB descriptor copy - and every kernel behind it - back for ~0.5 ms
    struct PullSeg { const void* src; void* dst; unsigned int bytes; unsigned int pad; };
    PullSeg* h_pull = nullptr;        // pinned + mapped segment lists
    int pull_used = 0, pull_cap = 0;
    SvoSeqCall* d_call = nullptr;
    SvoSeqCall* d_kcall = nullptr;    // svo_step_keyframe: descriptors of the keyframe stage
    int ring = 0;
    unsigned char* h_out = nullptr;   // pinned
    unsigned char* h_in = nullptr;    // pinned small-input staging
    size_t h_in_bytes = 0;
    unsigned char* h_trk = nullptr;   // staging of svo_tracker_set alone (that call does not wait)
    cudaEvent_t ev_trk = nullptr;     // recorded once the device has read h_trk

    std::vector<SeqState> seq;
    int64_t launches = 0;
    // optional per-kernel profiling (CUDA events on the engine stream)
    bool profiling = false;
    std::vector<cudaEvent_t> prof_ev;      // pool, pairs
    std::vector<int> prof_slot;            // slot id per pair in flight
    int prof_used = 0;
    double prof_ms[SVO_PROF_SLOTS] = {0};
    int64_t prof_count[SVO_PROF_SLOTS] = {0};
    cudaEvent_t timer_ev[2] = {nullptr, nullptr};
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    float last_ms[3] = {0, 0, 0};
};

namespace {

std::string g_create_err;

int fail(svo_engine* e, int code, const char* what, cudaError_t ce = cudaSuccess) {
    char buf[512];
    if (ce != cudaSuccess)
        snprintf(buf, sizeof buf, "%s: %s", what, cudaGetErrorString(ce));
    else
        snprintf(buf, sizeof buf, "%s", what);
    if (e) e->err = buf; else g_create_err = buf;
    return code;
}

#define ECUDA(e, expr)                                                        \
    do {                                                                      \
        cudaError_t _ce = (expr);                                             \
        if (_ce != cudaSuccess) return fail((e), SVO_ERR_CUDA, #expr, _ce);   \
    } while (0)

int round_up(int v, int m) { return (v + m - 1) / m * m; }

// profiling scope: records an event pair around a kernel group when profiling is on
struct ProfScope {
    svo_engine* e;
    int idx = -1;
    cudaStream_t st;
    ProfScope(svo_engine* eng, int slot, cudaStream_t stream = nullptr)
        : e(eng), st(stream ? stream : eng->st) {
        if (!e->profiling) return;
        if ((size_t)(e->prof_used + 1) * 2 > e->prof_ev.size()) {
            cudaEvent_t a, b;
            if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) return;
            e->prof_ev.push_back(a);
            e->prof_ev.push_back(b);
            e->prof_slot.push_back(0);
        }
        idx = e->prof_used++;
        e->prof_slot[idx] = slot;
        cudaEventRecord(e->prof_ev[2 * idx], st);
    }
    ~ProfScope() {
        if (idx >= 0) cudaEventRecord(e->prof_ev[2 * idx + 1], st);
    }
};

// after a stream synchronize: fold the finished event pairs into the per-kernel totals
void prof_collect(svo_engine* e) {
    for (int i = 0; i < e->prof_used; ++i) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, e->prof_ev[2 * i], e->prof_ev[2 * i + 1]) == cudaSuccess) {
            e->prof_ms[e->prof_slot[i]] += ms;
            e->prof_count[e->prof_slot[i]] += 1;
        }
    }
    e->prof_used = 0;
}

// ------------------------------------------------------------------ small glue kernels
__global__ void reset_kernel(unsigned int* eig_max, int* cand_count, int* n_corners,
                             const SvoSeqCall* call, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        int s = call[i].seq;
        eig_max[s] = 0u;
        cand_count[s] = 0;
        n_corners[s] = 0;
    }
}

// ingest: copy the (unpitched, arbitrarily aligned) source rows of the current stereo pair into
// the pitched device images (pyramid level 0 of the current slot, right image). A thread moves
// 16 output bytes: five aligned 32-bit loads, funnel shifts, one 128-bit store.
__global__ void __launch_bounds__(256)
ingest_kernel(uint8_t* pyr, SvoPyrLayout lay, size_t seq_stride, uint8_t* right,
              size_t right_stride, const SvoSeqCall* __restrict__ call) {
    const SvoSeqCall cd = call[blockIdx.z];
    const int W = lay.lvl[0].w, H = lay.lvl[0].h, dpitch = lay.lvl[0].pitch;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 16;
    if (y >= H || x0 >= W) return;
#pragma unroll
    for (int which = 0; which < 2; ++which) {
        const uint8_t* src = which == 0 ? cd.src_left : cd.src_right;
        if (!src) continue;
        uint8_t* dst = which == 0
            ? pyr + (size_t)cd.seq * seq_stride + (size_t)cd.cur_slot * lay.slot_bytes + lay.lvl[0].off
            : right + (size_t)cd.seq * right_stride;
        const uint8_t* p = src + (size_t)y * cd.src_pitch + x0;
        const int valid = min(16, W - x0);
        const unsigned mis = (unsigned)((uintptr_t)p & 3u);
        const unsigned int* pa = reinterpret_cast<const unsigned int*>(p - mis);
        const int nwords = (int)((mis + valid + 3) >> 2);      // aligned words holding valid bytes
        unsigned int w[5];
#pragma unroll
        for (int k = 0; k < 5; ++k) w[k] = (k < nwords) ? __ldcs(pa + k) : 0u;   // read once: evict first
        const unsigned sh = mis * 8;
        uint4 o;
        o.x = __funnelshift_r(w[0], w[1], sh);
        o.y = __funnelshift_r(w[1], w[2], sh);
        o.z = __funnelshift_r(w[2], w[3], sh);
        o.w = __funnelshift_r(w[3], w[4], sh);
        *reinterpret_cast<uint4*>(dst + (size_t)y * dpitch + x0) = o;
    }
}

// svo_keyframe_tracked: FeatureTracker::init(tracked inliers + new stereo-valid features) on the
// device. The tracked features already sit at the head of the tracker's feature buffer; the
// new ones are appended with ids first_id, first_id + 1, ... and the whole set becomes
// initial_features (feature_tracker.cpp:3-16, ids as bundle_adjuster.cpp:98-116 hands them out).
__global__ void __launch_bounds__(128)
tracker_append_kernel(float2* feat, unsigned long long* ids, float2* initial, int max_features,
                      const float2* __restrict__ valid2d, int pts_stride,
                      const int* __restrict__ n_valid, const SvoSeqCall* __restrict__ call) {
    const SvoSeqCall cd = call[blockIdx.x];
    if (cd.flags & SVO_CALL_SKIP) return;
    const int nt = cd.n_aux;
    const int nv = min(n_valid[cd.seq], max_features - nt);
    const size_t fb = ((size_t)cd.seq * 2 + cd.feat_buf) * max_features;
    const float2* v = valid2d + (size_t)cd.seq * pts_stride;
    for (int j = threadIdx.x; j < nt; j += blockDim.x) initial[fb + j] = feat[fb + j];
    for (int j = threadIdx.x; j < nv; j += blockDim.x) {
        const float2 p = v[j];
        feat[fb + nt + j] = p;
        initial[fb + nt + j] = p;
        ids[fb + nt + j] = cd.first_id + (unsigned long long)j;
    }
}

// svo_step_keyframe: the per-frame decision of ImageProcessor::process on the device
// (image_processor.cpp:23-30 "fewer than 4 corners", :27-52 first keyframe, :63-65 skip unless
// av_parallax > threshold or 40 % of the tracks are lost), and the descriptors the keyframe
// stage of the same call runs on: the tracker's feature buffer AFTER this frame's tracking and
// its kept count as the inlier list. Sequences that are no keyframe, or have no right image on
// the device yet, are marked SVO_CALL_SKIP.
__global__ void decide_kernel(const SvoSeqCall* __restrict__ call, SvoSeqCall* __restrict__ kcall,
                              const int* __restrict__ n_corners,
                              const SvoTrackStats* __restrict__ stats, float parallax_thresh,
                              int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    SvoSeqCall c = call[i];
    const SvoTrackStats st = stats[c.seq];
    const bool has_tracker = (c.flags & SVO_CALL_TRACKER) != 0;
    const bool tracked = has_tracker && st.n_attempted >= 0;
    int decision;
    if (n_corners[c.seq] < 4) decision = SVO_FRAME_TOO_FEW;
    else if (!has_tracker) decision = SVO_FRAME_FIRST;
    else if (st.av_parallax <= parallax_thresh && (double)st.percent_lost < 0.4) decision = SVO_FRAME_SKIP;
    else decision = SVO_FRAME_KEYFRAME;
    const bool key = decision == SVO_FRAME_FIRST || decision == SVO_FRAME_KEYFRAME;
    if (tracked) c.feat_buf = 1 - c.feat_buf;      // track_filter compacted into the other buffer
    c.n_aux = tracked ? st.n_kept : 0;
    c.n_feat = c.n_aux;
    c.n_pts = 0;
    c.decision = decision;
    if (!key || !(c.flags & SVO_CALL_HAS_RIGHT)) c.flags |= SVO_CALL_SKIP;
    kcall[i] = c;
}

// pack the results of one step into a contiguous block per listed sequence (one D2H copy)
__global__ void pack_step_kernel(unsigned char* out, size_t block, const float2* corners,
                                 const int* n_corners, const int* cand_count, int max_corners,
                                 const SvoTrackStats* stats, const float2* feat,
                                 const unsigned long long* ids, int max_features,
                                 const SvoSeqCall* call) {
    const SvoSeqCall cd = call[blockIdx.x];
    unsigned char* o = out + (size_t)blockIdx.x * block;
    StepHeader* h = reinterpret_cast<StepHeader*>(o);
    float2* oc = reinterpret_cast<float2*>(o + sizeof(StepHeader));
    float2* of = oc + max_corners;
    unsigned long long* oi = reinterpret_cast<unsigned long long*>(of + max_features);
    const int nc = n_corners[cd.seq];
    const SvoTrackStats s = stats[cd.seq];
    if (threadIdx.x == 0) {
        h->n_corners = nc;
        h->cand_count = cand_count[cd.seq];
        h->stats = s;
    }
    for (int i = threadIdx.x; i < nc; i += blockDim.x)
        oc[i] = corners[(size_t)cd.seq * max_corners + i];
    if (s.n_attempted >= 0 && (cd.flags & 1)) {
        // the kept set was compacted into the OTHER feature buffer
        const size_t fb = ((size_t)cd.seq * 2 + (1 - cd.feat_buf)) * max_features;
        for (int i = threadIdx.x; i < s.n_kept; i += blockDim.x) {
            of[i] = feat[fb + i];
            oi[i] = ids[fb + i];
        }
    }
}

__global__ void pack_key_kernel(unsigned char* out, size_t block, int cap, const float2* pts,
                                const int* n_new, const float2* valid2d, const float* pts3d,
                                const int* valid_idx, const short* disp16, const int* n_valid,
                                const SvoSeqCall* call) {
    const SvoSeqCall cd = call[blockIdx.x];
    unsigned char* o = out + (size_t)blockIdx.x * block;
    KeyHeader* h = reinterpret_cast<KeyHeader*>(o);
    float2* o_new = reinterpret_cast<float2*>(o + sizeof(KeyHeader));
    float2* o_v2 = o_new + cap;
    float* o_3d = reinterpret_cast<float*>(o_v2 + cap);
    int* o_vi = reinterpret_cast<int*>(o_3d + 3 * (size_t)cap);
    short* o_d = reinterpret_cast<short*>(o_vi + cap);
    const size_t b = (size_t)cd.seq * cap;
    const bool skip = (cd.flags & SVO_CALL_SKIP) != 0;
    const int nn = skip ? 0 : (n_new ? n_new[cd.seq] : cd.n_pts);
    const int nv = skip ? 0 : n_valid[cd.seq];
    if (threadIdx.x == 0) {
        h->n_new = nn; h->n_valid = nv; h->decision = cd.decision; h->n_tracked = cd.n_aux;
    }
    for (int i = threadIdx.x; i < nn; i += blockDim.x) {
        o_new[i] = pts[b + i];
        o_d[i] = disp16[b + i];
    }
    for (int i = threadIdx.x; i < nv; i += blockDim.x) {
        o_v2[i] = valid2d[b + i];
        o_vi[i] = valid_idx[b + i];
        o_3d[3 * i + 0] = pts3d[(b + i) * 3 + 0];
        o_3d[3 * i + 1] = pts3d[(b + i) * 3 + 1];
        o_3d[3 * i + 2] = pts3d[(b + i) * 3 + 2];
    }
}

SvoSeqCall* next_call_slot(svo_engine* e) {
    e->ring = (e->ring + 1) % CALL_RING;
    return e->h_call + (size_t)e->ring * e->n_seq;
}
SvoSeqCall* dev_call_slot(svo_engine* e) { return e->d_call + (size_t)e->ring * e->n_seq; }

int check_seq(svo_engine* e, int seq) {
    if (!e) return SVO_ERR_INVALID;
    if (seq < 0 || seq >= e->n_seq) return fail(e, SVO_ERR_INVALID, "sequence index out of range");
    return SVO_OK;
}

void fill_call(svo_engine* e, SvoSeqCall& c, int seq) {
    const SeqState& s = e->seq[seq];
    memset(&c, 0, sizeof c);
    c.seq = seq;
    c.cur_slot = s.cur_slot;
    c.feat_buf = s.feat_buf;
    c.n_feat = s.n_feat;
    c.n_initial = s.n_initial;
    c.flags = s.tracker_init ? 1 : 0;
    for (int i = 0; i < 4; ++i) c.pose[i * 5] = 1.f;
}

// choose the scratch slot for a new left frame (never FeatureTracker::last_image)
void begin_frame(svo_engine* e, int seq) {
    SeqState& s = e->seq[seq];
    s.cur_slot = (s.last_slot == 0) ? 1 : 0;
    s.have_frame = true;
    s.have_right = false;     // set again by the upload of this frame's right image
}

uint8_t* slot_ptr(svo_engine* e, int seq, int slot) {
    return e->d_pyr + (size_t)seq * e->seq_stride + (size_t)slot * e->lay.slot_bytes;
}

void prefetch_discard(svo_engine* e);

// H2D of one stereo pair into staging set `set` on stream `st` (one copy when the pair is
// contiguous in host memory, else one per image)
cudaError_t stage_pair(svo_engine* e, cudaStream_t st, int set, int seq, const uint8_t* left,
                       const uint8_t* right, int pitch) {
    const int W = e->p.width, H = e->p.height;
    uint8_t* base = e->d_stage + (size_t)set * e->stage_set + (size_t)seq * 2 * e->stage_img;
    if (pitch == W && right == left + (size_t)W * H)
        return cudaMemcpyAsync(base, left, (size_t)2 * W * H, cudaMemcpyHostToDevice, st);
    for (int which = 0; which < 2; ++which) {
        const uint8_t* img = which ? right : left;
        if (!img) continue;
        uint8_t* dst = base + (size_t)which * e->stage_img;
        cudaError_t ce = (pitch == W)
            ? cudaMemcpyAsync(dst, img, (size_t)W * H, cudaMemcpyHostToDevice, st)
            : cudaMemcpy2DAsync(dst, W, img, pitch, W, H, cudaMemcpyHostToDevice, st);
        if (ce != cudaSuccess) return ce;
    }
    return cudaSuccess;
}

// H2D of the stereo pairs of n sequences. Copies of single images pay a fixed cost each (a
// 466 KB KITTI image moves at 35 GB/s, a large copy at 55 GB/s), so left images that sit at a
// constant distance in host memory (per-camera ring buffers carved from one slab) and go to
// consecutive staging slots travel as ONE strided copy; everything else goes pair by pair.
cudaError_t stage_batch(svo_engine* e, cudaStream_t st, int set, int n, const int32_t* seqs,
                        const uint8_t* const* left, const uint8_t* const* right, int pitch) {
    const int W = e->p.width, H = e->p.height;
    const ptrdiff_t img = (ptrdiff_t)W * H;
    int i = 0;
    while (i < n) {
        int j = i + 1;
        if (pitch == W && j < n) {
            const ptrdiff_t stride = left[j] - left[i];
            if (stride >= img && stride < ((ptrdiff_t)1 << 30) && seqs[j] == seqs[i] + 1) {
                while (j + 1 < n && left[j + 1] - left[j] == stride && seqs[j + 1] == seqs[j] + 1) ++j;
                ++j;                                   // run [i, j)
            }
        }
        if (j - i >= 3) {
            uint8_t* base = e->d_stage + (size_t)set * e->stage_set + (size_t)seqs[i] * 2 * e->stage_img;
            cudaError_t ce = cudaMemcpy2DAsync(base, 2 * e->stage_img, left[i], (size_t)(left[i + 1] - left[i]),
                                               (size_t)img, (size_t)(j - i), cudaMemcpyHostToDevice, st);
            if (ce != cudaSuccess) return ce;
            for (int k = i; k < j; ++k) {
                const uint8_t* r = right ? right[k] : nullptr;
                if (!r) continue;
                ce = cudaMemcpyAsync(base + (size_t)(k - i) * 2 * e->stage_img + e->stage_img, r, (size_t)img,
                                     cudaMemcpyHostToDevice, st);
                if (ce != cudaSuccess) return ce;
            }
            i = j;
            continue;
        }
        cudaError_t ce = stage_pair(e, st, set, seqs[i], left[i], right ? right[i] : nullptr, pitch);
        if (ce != cudaSuccess) return ce;
        ++i;
    }
    return cudaSuccess;
}

// stage one image of the current stereo pair and record its device source in the descriptor
cudaError_t upload_image(svo_engine* e, SvoSeqCall& c, int which, const uint8_t* img, int pitch,
                         bool device_src) {
    const uint8_t* src = img;
    if (!device_src) {
        // never land in a staging set that holds announced (prefetched) frames
        if (e->pf_count == 2) prefetch_discard(e);
        if (e->pf_count == 1) e->cur_set = 1 - e->pf[e->pf_head].set;
        uint8_t* stage = e->d_stage + (size_t)e->cur_set * e->stage_set +
                         ((size_t)c.seq * 2 + which) * e->stage_img;
        const int W = e->p.width, H = e->p.height;
        cudaError_t ce = (pitch == W)
            ? cudaMemcpyAsync(stage, img, (size_t)W * H, cudaMemcpyHostToDevice, e->st)
            : cudaMemcpy2DAsync(stage, W, img, pitch, W, H, cudaMemcpyHostToDevice, e->st);
        if (ce != cudaSuccess) return ce;
        src = stage;
        pitch = W;
    }
    if (which == 0) c.src_left = src; else { c.src_right = src; e->seq[c.seq].have_right = true; }
    c.src_pitch = pitch;
    return cudaSuccess;
}
cudaError_t upload_left(svo_engine* e, SvoSeqCall& c, const uint8_t* img, int pitch, bool dev) {
    return upload_image(e, c, 0, img, pitch, dev);
}
cudaError_t upload_right(svo_engine* e, SvoSeqCall& c, const uint8_t* img, int pitch, bool dev) {
    return upload_image(e, c, 1, img, pitch, dev);
}

// descriptors -> device, then the ingest kernel moves the staged pairs into the pitched images
cudaError_t push_calls(svo_engine* e, int n);
struct PullList;
cudaError_t push_and_ingest(svo_engine* e, int n) {
    SVO_CUDA_OK(push_calls(e, n));
    dim3 blk(128, 2);
    dim3 grid((e->p.width + 2047) / 2048, (e->p.height + 1) / 2, n);
    ProfScope ps(e, P_INGEST);
    ingest_kernel<<<grid, blk, 0, e->st>>>(e->d_pyr, e->lay, e->seq_stride, e->d_right,
                                           e->right_stride, e->d_call + (size_t)e->ring * e->n_seq);
    e->launches += 1;
    return cudaGetLastError();
}

static_assert(sizeof(SvoSeqCall) % 8 == 0, "descriptors are pulled in 8-byte words");

// one block per segment: 8-byte words read from mapped host memory (ld.cv: never served from a
// stale cache line - the host rewrites these buffers between calls) and written to the device
__global__ void __launch_bounds__(128)
pull_kernel(const svo_engine::PullSeg* __restrict__ segs) {
    const svo_engine::PullSeg* sp = segs + blockIdx.x;
    const uint2* src = reinterpret_cast<const uint2*>(__ldcv(reinterpret_cast<const unsigned long long*>(&sp->src)));
    uint2* dst = reinterpret_cast<uint2*>(__ldcv(reinterpret_cast<const unsigned long long*>(&sp->dst)));
    const unsigned n = __ldcv(&sp->bytes) >> 3;
    for (unsigned i = threadIdx.x; i < n; i += blockDim.x) dst[i] = __ldcv(src + i);
}

// segment lists live in a ring of the pinned h_pull array (room for four lists); every API call
// but svo_tracker_set waits for the stream, so a list is consumed before its slots come round
struct PullList {
    svo_engine* e;
    int first, count;
    explicit PullList(svo_engine* eng) : e(eng), first(eng->pull_used), count(0) {
        if (first + 4 * eng->n_seq + 8 > eng->pull_cap) first = eng->pull_used = 0;
    }
    // src must lie in mapped pinned memory (h_call / h_in), bytes % 8 == 0
    void add(void* dst, const void* src, size_t bytes) {
        if (bytes == 0) return;
        svo_engine::PullSeg& sg = e->h_pull[first + count++];
        sg.src = src; sg.dst = dst; sg.bytes = (unsigned)bytes; sg.pad = 0;
    }
    cudaError_t launch() {
        e->pull_used = first + count;
        if (count == 0) return cudaSuccess;
        if (e->pf_count == 0) {
            // no bulk prefetch queued on the copy engine: plain DMA has the lower latency
            for (int i = 0; i < count; ++i) {
                const svo_engine::PullSeg& sg = e->h_pull[first + i];
                cudaError_t ce = cudaMemcpyAsync(sg.dst, sg.src, sg.bytes, cudaMemcpyHostToDevice, e->st);
                if (ce != cudaSuccess) return ce;
            }
            return cudaSuccess;
        }
        pull_kernel<<<count, 128, 0, e->st>>>(e->h_pull + first);
        e->launches += 1;
        return cudaGetLastError();
    }
};

// Per-call descriptors travel INSIDE a kernel launch (by-value parameter, 32 descriptors =
// 3.8 KB at a time) and are written to the device ring by that kernel: no DMA that could queue
// behind a bulk prefetch, and no read of host memory by an SM (slow while the link is busy).
struct CallChunk { SvoSeqCall c[32]; };
__global__ void __launch_bounds__(128)
store_calls_kernel(const __grid_constant__ CallChunk chunk, SvoSeqCall* __restrict__ dst, int n) {
    constexpr int WORDS = sizeof(SvoSeqCall) / 8;
    const unsigned long long* src = reinterpret_cast<const unsigned long long*>(chunk.c);
    unsigned long long* d = reinterpret_cast<unsigned long long*>(dst);
    for (int i = threadIdx.x; i < n * WORDS; i += blockDim.x) d[i] = src[i];
}

cudaError_t push_calls(svo_engine* e, int n) {
    const SvoSeqCall* hc = e->h_call + (size_t)e->ring * e->n_seq;
    SvoSeqCall* dc = dev_call_slot(e);
    for (int i0 = 0; i0 < n; i0 += 32) {
        CallChunk ch;
        const int m = std::min(32, n - i0);
        memcpy(ch.c, hc + i0, (size_t)m * sizeof(SvoSeqCall));
        store_calls_kernel<<<1, 128, 0, e->st>>>(ch, dc + i0, m);
        e->launches += 1;
    }
    return cudaGetLastError();
}

// levels 1.. of the current frames' pyramids (one launch per level, dependent launches)
cudaError_t launch_pyramid(svo_engine* e, const SvoSeqCall* dc, int n, cudaStream_t st) {
    // every build gets its own serial: lk_track keys what it caches about an image with it
    if (++e->pyr_serial == 0u) e->pyr_serial = 1u;
    return svo_launch_pyr_down(e->d_pyr, e->lay, e->seq_stride, dc, n, st, e->d_img_serial,
                               e->pyr_serial);
}

// pyramid + detection for the listed sequences (current frames already uploaded)
// lists: nullptr = the engine's per-sequence lists; the spill path passes its own
struct DetectLists {
    unsigned long long* strips;
    int strip_cap;
    unsigned long long* cand;
    int max_cand;
};
cudaError_t launch_detect(svo_engine* e, int n, bool with_pyramid, cudaStream_t st = nullptr,
                          const DetectLists* lists = nullptr) {
    if (!st) st = e->st;
    const SvoSeqCall* dc = dev_call_slot(e);
    const DetectLists own = {e->d_strips, e->strip_cap, e->d_cand, e->max_cand};
    const DetectLists& L = lists ? *lists : own;
    reset_kernel<<<(n + 127) / 128, 128, 0, st>>>(e->d_eig_max, e->d_cand_count, e->d_ncorners,
                                                      dc, n);
    e->launches += 1;
    if (with_pyramid) {
        ProfScope ps(e, P_PYR, st);
        SVO_CUDA_OK(launch_pyramid(e, dc, n, st));
        e->launches += e->lay.n_levels - 1;
    }
    e->cs_epoch = (e->cs_epoch % 0x3fffffffu) + 1u;    // tags this launch's verdict words
    const SvoCornerCall cc = {e->d_pyr, e->seq_stride, e->d_eig, e->eig_pitch, e->eig_stride,
                              e->d_eig_max, e->p.quality_level, L.strips, e->n_strips,
                              e->n_seg, L.strip_cap, e->d_strip_count, e->d_sub_max,
                              e->d_seg_flag, e->d_seg_state, e->d_cs_ticket, e->d_cs_work,
                              e->cs_epoch, dc};
    {
        ProfScope ps(e, P_SCORE, st);
        SVO_CUDA_OK(svo_launch_corner_score_nms(cc, e->lay, n, st));
    }
    {
        ProfScope ps(e, P_SORT, st);
        SVO_CUDA_OK(svo_launch_sort_candidates(L.cand, L.max_cand, e->d_cand_count,
                                               L.strips, e->n_strips * e->n_seg, L.strip_cap,
                                               e->d_strip_count, e->d_eig_max,
                                               e->p.quality_level, dc, n, st));
    }
    {
        ProfScope ps(e, P_GREEDY, st);
        SVO_CUDA_OK(svo_launch_greedy_select(L.cand, L.max_cand, e->d_cand_count, e->p.width,
                                             e->p.height, e->p.min_distance, e->p.max_corners,
                                             e->d_corners, e->d_ncorners, e->d_grid,
                                             e->grid_stride, dc, n, st));
    }
    e->launches += 2 + svo_sort_launch_count(L.max_cand);
    return cudaSuccess;
}

// the index map rebuilt by the last svo_tracker_set must be complete before anything on the
// engine stream reads or rewrites kept_idx
cudaError_t join_rekey(svo_engine* e) {
    if (!e->rekey_pending) return cudaSuccess;
    e->rekey_pending = false;
    return cudaStreamWaitEvent(e->st, e->ev_rekey, 0);
}

cudaError_t launch_track(svo_engine* e, int n, int max_feat_in_call, int flow_back, bool gated,
                         cudaEvent_t detect_done = nullptr) {
    const SvoSeqCall* dc = dev_call_slot(e);
    SvoLkParams lp;
    lp.max_iters = e->p.lk_max_iters;
    lp.eps2 = e->p.lk_epsilon * e->p.lk_epsilon;
    lp.min_eig = e->p.lk_min_eig_threshold;
    lp.grid_pts = max_feat_in_call;
    const int* gate = gated ? e->d_ncorners : nullptr;
    // detection still running on the other stream: the LK kernels cannot read the corner count
    // yet. They track unconditionally (scratch outputs only); the gate is applied by the filter.
    const int* gate_lk = detect_done ? nullptr : gate;
    const int mf = e->p.max_features;
    SvoLkCache lkc = {nullptr, nullptr, nullptr, nullptr, nullptr};
    SVO_CUDA_OK(join_rekey(e));
    if (e->lk_handover)
        lkc = SvoLkCache{e->d_lk_data, e->d_lk_hdr, e->d_img_serial, e->d_kept_idx,
                         e->lk_count ? e->d_lk_stats : nullptr};
    if (max_feat_in_call > 0) {
        {
            ProfScope ps(e, P_LKF);
            SVO_CUDA_OK(svo_launch_lk_track(e->d_pyr, e->lay, e->seq_stride, e->d_feat,
                                            (size_t)2 * mf, 1, e->d_tracked, e->d_status1, mf, 0,
                                            lp, mf, gate_lk, 4, nullptr, dc, n, e->st, lkc));
        }
        e->launches += 1;
        if (flow_back) {
            ProfScope ps(e, P_LKB);
            SVO_CUDA_OK(svo_launch_lk_track(e->d_pyr, e->lay, e->seq_stride, e->d_tracked,
                                            (size_t)mf, 0, e->d_reverse, e->d_status2, mf, 1, lp,
                                            mf, gate_lk, 4, e->d_status1, dc, n, e->st, lkc));
            e->launches += 1;
        }
    }
    if (detect_done) SVO_CUDA_OK(cudaStreamWaitEvent(e->st, detect_done, 0));
    ProfScope psf(e, P_FILTER);
    SVO_CUDA_OK(svo_launch_track_filter(e->d_feat, e->d_ids, e->d_initial, (size_t)2 * mf,
                                        e->d_tracked, e->d_status1, e->d_reverse, e->d_status2, mf,
                                        flow_back, e->p.fb_max_error, e->p.max_parallax,
                                        e->d_kept_idx, e->d_kept_par, e->d_stats, mf, gate, 4, dc,
                                        n, e->st));
    e->launches += 1;
    return cudaSuccess;
}

// Wait until everything queued on the engine stream is done. Spinning (the CUDA default) has the
// lowest latency; with more driver threads than host cores (8 ranks x 4 engine groups on a
// 32-core box) the spinning threads starve each other and the thread that should launch the
// next kernels: svo_set_blocking_sync() makes the wait sleep on an interrupt instead.
cudaError_t wait_stream(svo_engine* e) {
    if (!e->blocking_sync) return cudaStreamSynchronize(e->st);
    SVO_CUDA_OK(cudaEventRecord(e->ev_block, e->st));
    return cudaEventSynchronize(e->ev_block);
}

// Spill path of the detector. The engine's candidate lists are sized for ordinary images (a
// quarter of the pixels per segment, max_candidates survivors); an image with large plateaus
// of equal corner score can exceed either, and cv::goodFeaturesToTrack has no such limit. The
// detection of that ONE sequence is then repeated - its frame still sits in the sequence's
// current pyramid slot - on lists with room for every pixel, allocated at the first overflow.
// Leaves d_corners / d_ncorners / d_cand_count of the sequence exact; the stream is idle on
// return.
int detect_spill(svo_engine* e, int seq) {
    const size_t n_sub = (size_t)e->n_strips * e->n_seg;
    if (!e->d_strips_big) {
        e->strip_cap_big = svo_corner_strip_cap_full(e->p.height);
        e->max_cand_big = svo_sort_capacity_pad(e->p.width * e->p.height);
        cudaError_t ce = dalloc(&e->d_strips_big, n_sub * (size_t)e->strip_cap_big);
        if (ce == cudaSuccess) ce = dalloc(&e->d_cand_big, (size_t)e->max_cand_big);
        if (ce != cudaSuccess) {
            cudaGetLastError();
            if (e->d_strips_big) cudaFree(e->d_strips_big);
            e->d_strips_big = nullptr;
            return fail(e, SVO_ERR_NOMEM, "detector spill lists", ce);
        }
    }
    SvoSeqCall* hc = next_call_slot(e);
    fill_call(e, hc[0], seq);
    ECUDA(e, push_calls(e, 1));
    // the kernels index their lists by sequence: bias the single-sequence lists accordingly
    const uintptr_t strips0 = reinterpret_cast<uintptr_t>(e->d_strips_big) -
                              (uintptr_t)seq * n_sub * (size_t)e->strip_cap_big * 8u;
    const uintptr_t cand0 = reinterpret_cast<uintptr_t>(e->d_cand_big) -
                            (uintptr_t)seq * (size_t)e->max_cand_big * 8u;
    const DetectLists L = {reinterpret_cast<unsigned long long*>(strips0), e->strip_cap_big,
                           reinterpret_cast<unsigned long long*>(cand0), e->max_cand_big};
    ECUDA(e, launch_detect(e, 1, false, e->st, &L));
    ECUDA(e, wait_stream(e));
    e->spill_seq = seq;
    e->spills += 1;
    return SVO_OK;
}

int finish_timing(svo_engine* e) {
    float a = 0, b = 0, c = 0;
    cudaEventElapsedTime(&a, e->ev[0], e->ev[1]);
    cudaEventElapsedTime(&b, e->ev[1], e->ev[2]);
    cudaEventElapsedTime(&c, e->ev[2], e->ev[3]);
    e->last_ms[0] = a; e->last_ms[1] = b; e->last_ms[2] = c;
    if (e->profiling) prof_collect(e);
    return SVO_OK;
}

// drop every pending prefetch (the caller asked for different frames than it announced)
void prefetch_discard(svo_engine* e) {
    if (e->pf_count == 0) return;
    cudaStreamSynchronize(e->st_copy);
    for (auto& p : e->pf) p.valid = false;
    e->pf_count = 0;
    e->pf_head = 0;
}

// If the oldest pending prefetch announced exactly these frames: make the engine stream wait
// for its copies, point the descriptors at that staging set and return true.
bool prefetch_consume(svo_engine* e, int n, const int32_t* seqs, const uint8_t* const* left,
                      const uint8_t* const* right, int pitch, SvoSeqCall* hc) {
    if (e->pf_count == 0) return false;
    svo_engine::Prefetch& p = e->pf[e->pf_head];
    bool same = p.valid && p.n == n && p.pitch == pitch;
    for (int i = 0; same && i < n; ++i)
        same = p.seqs[i] == seqs[i] && p.left[i] == left[i] &&
               p.right[i] == (right ? right[i] : nullptr);
    if (!same) {
        // a step for other frames than announced: keep a single announced entry for the step
        // after this one (the usual "prefetch k+1, then step k" order of a first step), drop
        // everything if it is passed over twice or both staging sets are taken
        if (e->pf_count == 2 || ++p.skipped > 1) prefetch_discard(e);
        return false;
    }
    if (cudaStreamWaitEvent(e->st, p.done, 0) != cudaSuccess) {
        prefetch_discard(e);
        return false;
    }
    const int W = e->p.width;
    for (int i = 0; i < n; ++i) {
        uint8_t* base = e->d_stage + (size_t)p.set * e->stage_set +
                        (size_t)seqs[i] * 2 * e->stage_img;
        hc[i].src_left = base;
        hc[i].src_pitch = W;
        if (p.right[i]) {
            hc[i].src_right = base + e->stage_img;
            e->seq[seqs[i]].have_right = true;
        }
    }
    p.valid = false;
    e->cur_set = p.set;
    e->pf_head ^= 1;
    e->pf_count -= 1;
    return true;
}

// keyframe stage appended to a step (svo_step_keyframe)
struct StepKey {
    const float* poses;               // n x 16
    const uint64_t* first_new_id;     // n
    float parallax_thresh, focal, cx, cy, baseline;
    int32_t* decision;                // n, svo_frame_decision
    svo_float2* new_feats; int32_t* n_new; svo_float2* valid_2d; svo_float3* pts_3d;
    int32_t* valid_idx; int16_t* disp16; int32_t* n_valid;
};

int step_impl(svo_engine* e, int n, const int32_t* seqs, const uint8_t* const* left,
              const uint8_t* const* right, int pitch, bool device_src, svo_float2* corners,
              int32_t* n_corners, svo_track_stats* stats, svo_float2* tracked,
              uint64_t* tracked_ids, const StepKey* key = nullptr) {
    if (!e) return SVO_ERR_INVALID;
    if (n <= 0 || n > e->n_seq || !seqs || !left) return fail(e, SVO_ERR_INVALID, "svo_step: bad arguments");
    if (pitch < e->p.width) return fail(e, SVO_ERR_INVALID, "svo_step: pitch < width");
    ECUDA(e, cudaSetDevice(e->device));
    SvoSeqCall* hc = next_call_slot(e);
    int max_feat = 0;
    for (int i = 0; i < n; ++i) {
        int s = seqs[i];
        if (s < 0 || s >= e->n_seq) return fail(e, SVO_ERR_INVALID, "svo_step: sequence index out of range");
        begin_frame(e, s);
        fill_call(e, hc[i], s);
        if (e->seq[s].tracker_init && e->seq[s].n_feat > max_feat) max_feat = e->seq[s].n_feat;
    }
    ECUDA(e, cudaEventRecord(e->ev[0], e->st));
    if (device_src || !prefetch_consume(e, n, seqs, left, right, pitch, hc)) {
        if (!device_src) {
            // the staging set no pending prefetch occupies
            e->cur_set = e->pf_count ? 1 - e->pf[e->pf_head].set : 0;
        }
        if (!device_src) ECUDA(e, stage_batch(e, e->st, e->cur_set, n, seqs, left, right, pitch));
        for (int i = 0; i < n; ++i) {
            if (device_src) {
                ECUDA(e, upload_left(e, hc[i], left[i], pitch, true));
                if (right && right[i]) ECUDA(e, upload_right(e, hc[i], right[i], pitch, true));
                continue;
            }
            const uint8_t* r = right ? right[i] : nullptr;
            uint8_t* base = e->d_stage + (size_t)e->cur_set * e->stage_set +
                            (size_t)seqs[i] * 2 * e->stage_img;
            hc[i].src_left = base;
            hc[i].src_pitch = e->p.width;
            if (r) {
                hc[i].src_right = base + e->stage_img;
                e->seq[seqs[i]].have_right = true;
            }
        }
    }
    if (key) {
        for (int i = 0; i < n; ++i) {
            memcpy(hc[i].pose, key->poses + (size_t)i * 16, 16 * sizeof(float));
            hc[i].first_id = key->first_new_id[i];
            if (e->seq[seqs[i]].have_right) hc[i].flags |= SVO_CALL_HAS_RIGHT;
        }
    }
    ECUDA(e, cudaEventRecord(e->ev[1], e->st));
    ECUDA(e, push_and_ingest(e, n));
    if (e->split_streams && !e->profiling) {
        ECUDA(e, cudaEventRecord(e->ev_fork, e->st));
        ECUDA(e, cudaStreamWaitEvent(e->st_det, e->ev_fork, 0));
        ECUDA(e, launch_detect(e, n, false, e->st_det));
        ECUDA(e, cudaEventRecord(e->ev_join, e->st_det));
        ECUDA(e, launch_pyramid(e, dev_call_slot(e), n, e->st));
        e->launches += e->lay.n_levels - 1;
        ECUDA(e, launch_track(e, n, max_feat, 1, true, e->ev_join));
    } else {
        ECUDA(e, launch_detect(e, n, true));
        ECUDA(e, launch_track(e, n, max_feat, 1, true));
    }
    {
        ProfScope ps(e, P_PACK);
        pack_step_kernel<<<n, 128, 0, e->st>>>(e->d_out, e->step_block, e->d_corners,
                                               e->d_ncorners, e->d_cand_count, e->p.max_corners,
                                               e->d_stats, e->d_feat, e->d_ids, e->p.max_features,
                                               dev_call_slot(e));
    }
    e->launches += 1;
    ECUDA(e, cudaGetLastError());
    size_t out_bytes = (size_t)n * e->step_block;
    if (key) {
        // decision + keyframe stage for the sequences that need (and can have) one, on the
        // descriptors the decision kernel derives; results go behind the step blocks
        const int mf = e->p.max_features, cap = e->pts_cap, mcn = e->p.max_corners;
        SvoSeqCall* kc = e->d_kcall;
        decide_kernel<<<(n + 63) / 64, 64, 0, e->st>>>(dev_call_slot(e), kc, e->d_ncorners,
                                                       e->d_stats, key->parallax_thresh, n);
        {
            ProfScope ps(e, P_DEDUPE);
            ECUDA(e, svo_launch_dedupe(e->d_corners, e->d_ncorners, mcn, e->d_feat, 2 * mf, mf,
                                       e->p.min_distance, e->d_pts, e->d_keep_idx, e->d_nnew, cap,
                                       kc, e->d_ticket, n, e->st));
        }
        SvoStereoParams sp = {e->p.num_disparities, e->p.sad_window, e->p.prefilter_cap,
                              e->p.texture_threshold, e->p.uniqueness_ratio};
        {
            ProfScope ps(e, P_STEREO);
            ECUDA(e, svo_launch_stereo_bm_sparse(e->d_pyr, e->lay, e->seq_stride, e->d_right,
                                                 e->right_stride, e->d_pts, cap, e->d_disp16, sp,
                                                 mcn, kc, n, e->st));
        }
        {
            ProfScope ps(e, P_REPROJ);
            ECUDA(e, svo_launch_reproject(e->d_pts, e->d_disp16, cap, key->focal, key->cx, key->cy,
                                          key->baseline, e->d_valid2d, e->d_pts3d, e->d_valid_idx,
                                          e->d_nvalid, kc, n, e->st));
        }
        tracker_append_kernel<<<n, 128, 0, e->st>>>(e->d_feat, e->d_ids, e->d_initial, mf,
                                                    e->d_valid2d, cap, e->d_nvalid, kc);
        pack_key_kernel<<<n, 128, 0, e->st>>>(e->d_out + out_bytes, e->key_block, cap, e->d_pts,
                                              e->d_nnew, e->d_valid2d, e->d_pts3d, e->d_valid_idx,
                                              e->d_disp16, e->d_nvalid, kc);
        e->launches += 6;
        ECUDA(e, cudaGetLastError());
        out_bytes += (size_t)n * e->key_block;
    }
    ECUDA(e, cudaEventRecord(e->ev[2], e->st));
    ECUDA(e, cudaMemcpyAsync(e->h_out, e->d_out, out_bytes, cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaEventRecord(e->ev[3], e->st));
    ECUDA(e, wait_stream(e));
    finish_timing(e);
    int rc = SVO_OK;
    const int mc = e->p.max_corners, mf = e->p.max_features;
    for (int i = 0; i < n; ++i) {
        // candidate lists overflowed (plateau image): detect this sequence again on the spill
        // lists and patch its block. Not possible once the keyframe stage of the same call has
        // consumed the corners (svo_step_keyframe): that call reports the overflow.
        unsigned char* o = e->h_out + (size_t)i * e->step_block;
        StepHeader* h = reinterpret_cast<StepHeader*>(o);
        if (h->cand_count <= e->max_cand) continue;
        if (key) {
            rc = fail(e, SVO_ERR_CAPACITY, "svo_step_keyframe: NMS candidate lists overflowed (plateau image); "
                                           "use svo_step + svo_keyframe[_tracked] for such frames or raise max_candidates");
            continue;
        }
        const int before = h->n_corners;
        const int src = detect_spill(e, seqs[i]);
        if (src != SVO_OK) { rc = src; continue; }
        int hdr[2] = {0, 0};
        ECUDA(e, cudaMemcpyAsync(&hdr[0], e->d_ncorners + seqs[i], sizeof(int), cudaMemcpyDeviceToHost, e->st));
        ECUDA(e, cudaMemcpyAsync(&hdr[1], e->d_cand_count + seqs[i], sizeof(int), cudaMemcpyDeviceToHost, e->st));
        ECUDA(e, cudaMemcpyAsync(o + sizeof(StepHeader), e->d_corners + (size_t)seqs[i] * mc,
                                 (size_t)mc * sizeof(float2), cudaMemcpyDeviceToHost, e->st));
        ECUDA(e, wait_stream(e));
        h->n_corners = hdr[0];
        h->cand_count = hdr[1];
        if (hdr[1] > e->max_cand_big)
            rc = fail(e, SVO_ERR_CAPACITY, "detector spill lists overflowed (internal error)");
        else if ((before >= 4) != (hdr[0] >= 4))
            // the device-side "fewer than 4 corners" gate of the tracking stage saw the
            // truncated list
            rc = fail(e, SVO_ERR_CAPACITY, "NMS candidate lists overflowed on a frame with fewer than 4 corners: raise max_candidates");
    }
    for (int i = 0; i < n; ++i) {
        const unsigned char* o = e->h_out + (size_t)i * e->step_block;
        const StepHeader* h = reinterpret_cast<const StepHeader*>(o);
        const float2* oc = reinterpret_cast<const float2*>(o + sizeof(StepHeader));
        const float2* of = oc + mc;
        const unsigned long long* oi = reinterpret_cast<const unsigned long long*>(of + mf);
        SeqState& s = e->seq[seqs[i]];
        if (n_corners) n_corners[i] = h->n_corners;
        if (corners) memcpy(corners + (size_t)i * mc, oc, (size_t)h->n_corners * sizeof(float2));
        svo_track_stats st = {0.f, 0.f, 0, 0};
        if (s.tracker_init && h->stats.n_attempted >= 0) {
            // FeatureTracker::track_features ran: the tracker now holds the kept set and this
            // frame as last_image (feature_tracker.cpp:39-41,66)
            st.av_parallax = h->stats.av_parallax;
            st.percent_lost = h->stats.percent_lost;
            st.n_kept = h->stats.n_kept;
            st.n_attempted = h->stats.n_attempted;
            s.dbg_attempted = h->stats.n_attempted;
            s.feat_buf = 1 - s.feat_buf;
            s.n_feat = h->stats.n_kept;
            s.last_slot = s.cur_slot;
            if (tracked) memcpy(tracked + (size_t)i * mf, of, (size_t)st.n_kept * sizeof(float2));
            if (tracked_ids) memcpy(tracked_ids + (size_t)i * mf, oi, (size_t)st.n_kept * sizeof(uint64_t));
        }
        if (stats) stats[i] = st;
    }
    if (key) {
        const int cap = e->pts_cap, mcn = e->p.max_corners;
        const unsigned char* kb = e->h_out + (size_t)n * e->step_block;
        for (int i = 0; i < n; ++i) {
            const unsigned char* o = kb + (size_t)i * e->key_block;
            const KeyHeader* h = reinterpret_cast<const KeyHeader*>(o);
            const float2* o_new = reinterpret_cast<const float2*>(o + sizeof(KeyHeader));
            const float2* o_v2 = o_new + cap;
            const float* o_3d = reinterpret_cast<const float*>(o_v2 + cap);
            const int* o_vi = reinterpret_cast<const int*>(o_3d + 3 * (size_t)cap);
            const short* o_d = reinterpret_cast<const short*>(o_vi + cap);
            const size_t ob = (size_t)i * mcn;
            SeqState& st = e->seq[seqs[i]];
            const bool is_key = h->decision == SVO_FRAME_FIRST || h->decision == SVO_FRAME_KEYFRAME;
            const bool done = is_key && st.have_right;
            // a keyframe without its right image on the device is reported as SVO_FRAME_*_PENDING:
            // the caller supplies the image (svo_set_right) and finishes with svo_keyframe_tracked
            key->decision[i] = (is_key && !done) ? h->decision + SVO_FRAME_PENDING : h->decision;
            key->n_new[i] = done ? h->n_new : 0;
            key->n_valid[i] = done ? h->n_valid : 0;
            if (!done) continue;
            if (key->new_feats) memcpy(key->new_feats + ob, o_new, (size_t)h->n_new * sizeof(float2));
            if (key->disp16) memcpy(key->disp16 + ob, o_d, (size_t)h->n_new * sizeof(short));
            if (key->valid_2d) memcpy(key->valid_2d + ob, o_v2, (size_t)h->n_valid * sizeof(float2));
            if (key->valid_idx) memcpy(key->valid_idx + ob, o_vi, (size_t)h->n_valid * sizeof(int));
            if (key->pts_3d) memcpy(key->pts_3d + ob, o_3d, (size_t)h->n_valid * 3 * sizeof(float));
            // the host mirror of what tracker_append_kernel did (FeatureTracker::init)
            const int nt = h->n_tracked;
            st.n_feat = nt + std::min((int)h->n_valid, mf - nt);
            st.n_initial = st.n_feat;
            st.tracker_init = true;
            st.last_slot = st.cur_slot;
        }
    }
    return rc;
}

}  // namespace

// =============================================================================== C ABI
extern "C" {

void svo_default_params(svo_params* p, int32_t width, int32_t height) {
    if (!p) return;
    memset(p, 0, sizeof *p);
    p->width = width;
    p->height = height;
    p->max_corners = 300;
    p->quality_level = 0.1;
    p->min_distance = 30.f;
    p->max_candidates = 0;
    p->num_disparities = 48;
    p->sad_window = 21;
    p->prefilter_cap = 31;
    p->texture_threshold = 10;
    p->uniqueness_ratio = 15;
    p->lk_max_level = 3;
    p->lk_max_iters = 30;
    p->lk_epsilon = 0.01;
    p->lk_min_eig_threshold = 1e-2;
    p->fb_max_error = 2.f;
    p->max_parallax = 200.f;
    p->max_features = 400;
    p->keep_eig_map = 0;
}

const char* svo_last_error(const svo_engine* e) { return e ? e->err.c_str() : g_create_err.c_str(); }

int svo_create(const svo_params* pp, int32_t device, int32_t n_seq, svo_engine** out) {
    if (!pp || !out || n_seq <= 0) return fail(nullptr, SVO_ERR_INVALID, "svo_create: bad arguments");
    *out = nullptr;
    svo_params p = *pp;
    if (p.width < 64 || p.height < 32 || p.width > 16384 || p.height > 16384)
        return fail(nullptr, SVO_ERR_INVALID, "svo_create: image size out of range (64x32 .. 16384x16384)");
    if (p.max_corners <= 0 || p.max_features <= 0 || p.max_features < 1)
        return fail(nullptr, SVO_ERR_INVALID, "svo_create: max_corners / max_features must be positive");
    if (p.lk_max_level < 0 || p.lk_max_level > 3)      // lk_track pre-stages <= 4 levels
        return fail(nullptr, SVO_ERR_INVALID, "svo_create: lk_max_level out of range");
    if (p.num_disparities <= 0 || p.num_disparities % 16 || p.sad_window < 5 || !(p.sad_window & 1) ||
        p.sad_window > 31 || p.num_disparities > 256)
        return fail(nullptr, SVO_ERR_INVALID, "svo_create: StereoBM parameters out of range");
    if ((long long)p.width * p.height >= (1ll << 31))
        return fail(nullptr, SVO_ERR_INVALID, "svo_create: image too large");

    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0)
        return fail(nullptr, SVO_ERR_CUDA, "svo_create: no CUDA device (this library has no CPU path)", ce);
    if (device < 0 || device >= ndev) return fail(nullptr, SVO_ERR_INVALID, "svo_create: device index out of range");
    ce = cudaSetDevice(device);
    if (ce != cudaSuccess) return fail(nullptr, SVO_ERR_CUDA, "cudaSetDevice", ce);

    svo_engine* e = new svo_engine();
    e->p = p;
    e->device = device;
    e->n_seq = n_seq;
    e->seq.resize(n_seq);

    // pyramid layout: levels as cv::buildOpticalFlowPyramid keeps them (a level must stay
    // larger than the 21x21 window)
    SvoPyrLayout& lay = e->lay;
    memset(&lay, 0, sizeof lay);
    int w = p.width, h = p.height;
    unsigned int off = 0;
    for (int l = 0; l <= p.lk_max_level; ++l) {
        if (l > 0) {
            int nw = (w + 1) / 2, nh = (h + 1) / 2;
            if (nw <= SVO_LK_WIN || nh <= SVO_LK_WIN) break;
            w = nw; h = nh;
        }
        lay.lvl[l].w = w;
        lay.lvl[l].h = h;
        lay.lvl[l].pitch = round_up(w, 128);
        lay.lvl[l].off = off;
        off += (unsigned int)lay.lvl[l].pitch * h;
        off = (off + 255u) & ~255u;
        lay.n_levels = l + 1;
    }
    lay.slot_bytes = off;
    e->seq_stride = (size_t)2 * off;
    e->pitch0 = lay.lvl[0].pitch;
    e->right_stride = ((size_t)e->pitch0 * p.height + 255) & ~(size_t)255;
    e->eig_pitch = round_up(p.width, 32);
    e->eig_stride = (size_t)e->eig_pitch * p.height;
    long long npx = (long long)p.width * p.height;
    int mc = p.max_candidates;
    if (mc <= 0) mc = npx <= (1 << 20) ? 16384 : (int)((npx / 16 < (1 << 15)) ? (1 << 15) : npx / 16);
    if (mc > 16384) mc = svo_sort_capacity_pad(mc);
    e->max_cand = mc;
    e->p.max_candidates = mc;
    e->grid_stride = (size_t)svo_greedy_grid_ints(p.width, p.height, p.min_distance) + 8;
    e->pts_cap = p.max_corners > p.max_features ? p.max_corners : p.max_features;
    const int cap = e->pts_cap, mf = p.max_features, mcn = p.max_corners;
    e->step_block = (sizeof(StepHeader) + (size_t)mcn * 8 + (size_t)mf * 16 + 15) & ~(size_t)15;
    e->key_block = (sizeof(KeyHeader) + (size_t)cap * (8 + 8 + 12 + 4 + 2) + 15) & ~(size_t)15;
    e->out_bytes = (e->step_block + e->key_block) * n_seq;   // svo_step_keyframe returns both
    e->h_in_bytes = (size_t)n_seq * ((size_t)cap * 16 + (size_t)mf * 16) + 4096;

#define CCHK(expr)                                                               \
    do {                                                                         \
        cudaError_t _ce = (expr);                                                \
        if (_ce != cudaSuccess) {                                                \
            fail(nullptr, _ce == cudaErrorMemoryAllocation ? SVO_ERR_NOMEM : SVO_ERR_CUDA, #expr, _ce); \
            svo_destroy(e);                                                      \
            return _ce == cudaErrorMemoryAllocation ? SVO_ERR_NOMEM : SVO_ERR_CUDA; \
        }                                                                        \
    } while (0)

    CCHK(cudaStreamCreateWithFlags(&e->st, cudaStreamNonBlocking));
    CCHK(cudaStreamCreateWithFlags(&e->st_copy, cudaStreamNonBlocking));
    CCHK(cudaStreamCreateWithFlags(&e->st_det, cudaStreamNonBlocking));
    CCHK(cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
    CCHK(cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming));
    CCHK(cudaEventCreateWithFlags(&e->ev_block, cudaEventDisableTiming | cudaEventBlockingSync));
    if (const char* v = getenv("SVO_SPLIT_STREAMS")) e->split_streams = atoi(v) != 0;
    for (auto& pfe : e->pf) CCHK(cudaEventCreateWithFlags(&pfe.done, cudaEventDisableTiming));
    for (int i = 0; i < 4; ++i) CCHK(cudaEventCreate(&e->ev[i]));
    for (int i = 0; i < 2; ++i) CCHK(cudaEventCreate(&e->timer_ev[i]));
    CCHK(svo_sort_init());
    CCHK(svo_greedy_init());
    const size_t ns = (size_t)n_seq;
    CCHK(dalloc(&e->d_pyr, ns * e->seq_stride));
    CCHK(dalloc(&e->d_right, ns * e->right_stride));
    // exact stride: a host pair stored back to back (right == left + W*H) lands with ONE copy;
    // the ingest kernel reads the staged images with aligned words + funnel shifts
    e->stage_img = (size_t)p.width * p.height;
    e->stage_set = ns * 2 * e->stage_img;
    CCHK(dalloc(&e->d_stage, 2 * e->stage_set));
    if (p.keep_eig_map) CCHK(dalloc(&e->d_eig, ns * e->eig_stride));
    e->n_strips = svo_corner_strips(p.width);
    e->n_seg = svo_corner_segments(p.height);
    e->strip_cap = svo_corner_strip_cap(p.height);
    const size_t n_sub = ns * e->n_strips * e->n_seg;
    CCHK(dalloc(&e->d_strips, n_sub * (size_t)e->strip_cap));
    CCHK(dalloc(&e->d_strip_count, n_sub));
    CCHK(dalloc(&e->d_sub_max, n_sub));
    CCHK(dalloc(&e->d_seg_flag, 2 * n_sub));
    CCHK(dalloc(&e->d_cs_work, 2));
    CCHK(cudaMemsetAsync(e->d_cs_work, 0, 2 * sizeof(unsigned int), e->st));
    CCHK(dalloc(&e->d_seg_state, n_sub * svo_corner_state_doubles()));
    CCHK(dalloc(&e->d_cs_ticket, ns));
    CCHK(cudaMemsetAsync(e->d_cs_ticket, 0, ns * sizeof(unsigned int), e->st));
    CCHK(cudaMemsetAsync(e->d_seg_flag, 0, 2 * n_sub * sizeof(unsigned int), e->st));
    CCHK(dalloc(&e->d_eig_max, ns));
    CCHK(dalloc(&e->d_cand, ns * mc));
    CCHK(dalloc(&e->d_cand_count, ns));
    CCHK(dalloc(&e->d_corners, ns * mcn));
    CCHK(dalloc(&e->d_ncorners, ns));
    CCHK(dalloc(&e->d_grid, ns * e->grid_stride));
    CCHK(dalloc(&e->d_feat, ns * 2 * mf));
    CCHK(dalloc(&e->d_ids, ns * 2 * mf));
    CCHK(dalloc(&e->d_initial, ns * 2 * mf));
    CCHK(dalloc(&e->d_tracked, ns * mf));
    CCHK(dalloc(&e->d_status1, ns * mf));
    CCHK(dalloc(&e->d_reverse, ns * mf));
    CCHK(dalloc(&e->d_status2, ns * mf));
    CCHK(dalloc(&e->d_kept_idx, ns * mf));
    CCHK(dalloc(&e->d_kept_par, ns * mf));
    CCHK(cudaMemsetAsync(e->d_kept_idx, 0, ns * mf * sizeof(int), e->st));
    CCHK(dalloc(&e->d_img_serial, ns * 2));
    CCHK(cudaMemsetAsync(e->d_img_serial, 0, ns * 2 * sizeof(unsigned), e->st));
    {
        // 12 KB per tracked feature (4 levels x 63 window threads x 24 int16); SVO_LK_HANDOVER=0
        // switches the hand-over off, and it is left out when it would not fit comfortably
        static const bool want = [] { const char* v = getenv("SVO_LK_HANDOVER"); return !v || atoi(v) != 0; }();
        const size_t data_bytes = ns * mf * (size_t)4 * 3 * 64 * sizeof(uint4);
        if (want && e->lay.n_levels >= 2 && data_bytes <= ((size_t)8 << 30)) {
            CCHK(dalloc(&e->d_lk_data, data_bytes / sizeof(uint4)));
            CCHK(dalloc(&e->d_lk_hdr, ns * mf));
            CCHK(cudaMemsetAsync(e->d_lk_hdr, 0, ns * mf * sizeof(SvoLkCacheHdr), e->st));
            CCHK(dalloc(&e->d_lk_stats, 2));
            CCHK(cudaMemsetAsync(e->d_lk_stats, 0, 2 * sizeof(unsigned long long), e->st));
            e->lk_handover = true;
            if (const char* v = getenv("SVO_LK_REKEY")) e->lk_rekey = atoi(v) != 0;
            if (e->lk_rekey) {
                int least = 0, greatest = 0;
                CCHK(cudaDeviceGetStreamPriorityRange(&least, &greatest));
                CCHK(cudaStreamCreateWithPriority(&e->st_rekey, cudaStreamNonBlocking, greatest));
                CCHK(cudaEventCreateWithFlags(&e->ev_feat, cudaEventDisableTiming));
                CCHK(cudaEventCreateWithFlags(&e->ev_rekey, cudaEventDisableTiming));
            }
        }
    }
    CCHK(dalloc(&e->d_stats, ns));
    CCHK(dalloc(&e->d_inl, ns * mf));
    CCHK(dalloc(&e->d_pts, ns * cap));
    CCHK(dalloc(&e->d_keep_idx, ns * cap));
    CCHK(dalloc(&e->d_nnew, ns));
    CCHK(dalloc(&e->d_ticket, ns));
    CCHK(cudaMemsetAsync(e->d_ticket, 0, ns * sizeof(unsigned), e->st));
    CCHK(dalloc(&e->d_disp16, ns * cap));
    CCHK(dalloc(&e->d_valid2d, ns * cap));
    CCHK(dalloc(&e->d_pts3d, ns * cap * 3));
    CCHK(dalloc(&e->d_valid_idx, ns * cap));
    CCHK(dalloc(&e->d_nvalid, ns));
    CCHK(dalloc(&e->d_out, e->out_bytes));
    CCHK(dalloc(&e->d_call, (size_t)CALL_RING * ns));
    CCHK(dalloc(&e->d_kcall, ns));
    // cudaHostAllocMapped: with unified addressing the device uses the same pointers
    CCHK(cudaHostAlloc(reinterpret_cast<void**>(&e->h_call), (size_t)CALL_RING * ns * sizeof(SvoSeqCall),
                       cudaHostAllocMapped | cudaHostAllocPortable));
    e->pull_cap = 16 * n_seq + 64;
    CCHK(cudaHostAlloc(reinterpret_cast<void**>(&e->h_pull), (size_t)e->pull_cap * sizeof(svo_engine::PullSeg),
                       cudaHostAllocMapped | cudaHostAllocPortable));
    CCHK(cudaMallocHost(reinterpret_cast<void**>(&e->h_out), e->out_bytes));
    CCHK(cudaHostAlloc(reinterpret_cast<void**>(&e->h_in), e->h_in_bytes,
                       cudaHostAllocMapped | cudaHostAllocPortable));
    CCHK(cudaHostAlloc(reinterpret_cast<void**>(&e->h_trk), ns * mf * 16 + 64,
                       cudaHostAllocMapped | cudaHostAllocPortable));
    CCHK(cudaEventCreateWithFlags(&e->ev_trk, cudaEventDisableTiming));
    CCHK(cudaMemsetAsync(e->d_pyr, 0, ns * e->seq_stride, e->st));
    CCHK(cudaMemsetAsync(e->d_right, 0, ns * e->right_stride, e->st));
    if (e->d_eig) CCHK(cudaMemsetAsync(e->d_eig, 0, ns * e->eig_stride * sizeof(float), e->st));
    CCHK(cudaMemsetAsync(e->d_status2, 0, ns * mf, e->st));
    CCHK(cudaMemsetAsync(e->d_reverse, 0, ns * mf * sizeof(float2), e->st));
    CCHK(cudaMemsetAsync(e->d_stats, 0, ns * sizeof(SvoTrackStats), e->st));
    CCHK(cudaStreamSynchronize(e->st));
#undef CCHK
    *out = e;
    return SVO_OK;
}

void svo_destroy(svo_engine* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->st) cudaStreamSynchronize(e->st);
    if (e->st_copy) cudaStreamSynchronize(e->st_copy);
    if (e->st_det) cudaStreamSynchronize(e->st_det);
    if (e->st_rekey) cudaStreamSynchronize(e->st_rekey);
    void* dptrs[] = {e->d_stage, e->d_pyr, e->d_right, e->d_dense, e->d_eig, e->d_eig_max, e->d_cand, e->d_cand_count,
                     e->d_strips, e->d_strip_count, e->d_sub_max, e->d_seg_flag, e->d_seg_state, e->d_cs_ticket, e->d_cs_work, e->d_corners, e->d_ncorners, e->d_grid, e->d_feat, e->d_ids, e->d_initial,
                     e->d_tracked, e->d_status1, e->d_reverse, e->d_status2, e->d_kept_idx,
                     e->d_kept_par, e->d_stats, e->d_inl, e->d_pts, e->d_keep_idx, e->d_nnew, e->d_ticket,
                     e->d_disp16, e->d_valid2d, e->d_pts3d, e->d_valid_idx, e->d_nvalid, e->d_out,
                     e->d_call, e->d_kcall, e->d_pnp_world, e->d_pnp_img, e->d_pnp_counts,
                     e->d_pnp_guess, e->d_pnp_res, e->d_pnp_inl, e->d_ba, e->d_draw,
                     e->d_strips_big, e->d_cand_big, e->d_lk_data, e->d_lk_hdr, e->d_img_serial, e->d_lk_stats};
    for (void* p : dptrs) if (p) cudaFree(p);
    if (e->h_call) cudaFreeHost(e->h_call);
    if (e->h_pull) cudaFreeHost(e->h_pull);
    if (e->h_out) cudaFreeHost(e->h_out);
    if (e->h_in) cudaFreeHost(e->h_in);
    if (e->h_trk) cudaFreeHost(e->h_trk);
    if (e->ev_trk) cudaEventDestroy(e->ev_trk);
    for (int i = 0; i < 4; ++i) if (e->ev[i]) cudaEventDestroy(e->ev[i]);
    for (int i = 0; i < 2; ++i) if (e->timer_ev[i]) cudaEventDestroy(e->timer_ev[i]);
    for (cudaEvent_t ev : e->prof_ev) cudaEventDestroy(ev);
    for (auto& pfe : e->pf) if (pfe.done) cudaEventDestroy(pfe.done);
    if (e->st_copy) cudaStreamDestroy(e->st_copy);
    if (e->st_det) cudaStreamDestroy(e->st_det);
    if (e->st_rekey) cudaStreamDestroy(e->st_rekey);
    if (e->ev_feat) cudaEventDestroy(e->ev_feat);
    if (e->ev_rekey) cudaEventDestroy(e->ev_rekey);
    if (e->ev_fork) cudaEventDestroy(e->ev_fork);
    if (e->ev_join) cudaEventDestroy(e->ev_join);
    if (e->ev_block) cudaEventDestroy(e->ev_block);
    if (e->st) cudaStreamDestroy(e->st);
    delete e;
}

int svo_get_params(const svo_engine* e, svo_params* out) {
    if (!e || !out) return SVO_ERR_INVALID;
    *out = e->p;
    return SVO_OK;
}
int svo_num_levels(const svo_engine* e) { return e ? e->lay.n_levels : SVO_ERR_INVALID; }
int64_t svo_detector_spills(const svo_engine* e) { return e ? e->spills : 0; }
int svo_lk_handover(svo_engine* e, int enable, int count, int64_t* lookups, int64_t* hits) {
    if (!e) return SVO_ERR_INVALID;
    cudaSetDevice(e->device);
    if (enable == 1 && !e->d_lk_data)
        return fail(e, SVO_ERR_INVALID, "svo_lk_handover: this engine has no hand-over buffers");
    if (enable >= 0) e->lk_handover = enable != 0;
    if (count >= 0) e->lk_count = count != 0 && e->d_lk_stats;
    unsigned long long st[2] = {0, 0};
    if ((lookups || hits) && e->d_lk_stats) {
        ECUDA(e, cudaMemcpyAsync(st, e->d_lk_stats, sizeof(st), cudaMemcpyDeviceToHost, e->st));
        ECUDA(e, cudaStreamSynchronize(e->st));
    }
    if (lookups) *lookups = (int64_t)st[0];
    if (hits) *hits = (int64_t)st[1];
    return SVO_OK;
}
int64_t svo_kernel_launches(const svo_engine* e) { return e ? e->launches : 0; }
float svo_last_gpu_ms(const svo_engine* e) { return e ? e->last_ms[1] : 0.f; }
int svo_last_timing(const svo_engine* e, float ms[3]) {
    if (!e || !ms) return SVO_ERR_INVALID;
    ms[0] = e->last_ms[0]; ms[1] = e->last_ms[1]; ms[2] = e->last_ms[2];
    return SVO_OK;
}

int svo_set_blocking_sync(svo_engine* e, int32_t on) {
    if (!e) return SVO_ERR_INVALID;
    e->blocking_sync = on != 0;
    return SVO_OK;
}

int svo_set_profiling(svo_engine* e, int32_t on) {
    if (!e) return SVO_ERR_INVALID;
    cudaStreamSynchronize(e->st);
    e->profiling = on != 0;
    e->prof_used = 0;
    for (int i = 0; i < SVO_PROF_SLOTS; ++i) { e->prof_ms[i] = 0; e->prof_count[i] = 0; }
    return SVO_OK;
}
int svo_profile_slots(void) { return SVO_PROF_SLOTS; }
const char* svo_profile_name(int32_t slot) {
    return (slot >= 0 && slot < SVO_PROF_SLOTS) ? kProfNames[slot] : "";
}
int svo_get_profile(const svo_engine* e, double* ms, int64_t* counts) {
    if (!e || !ms || !counts) return SVO_ERR_INVALID;
    for (int i = 0; i < SVO_PROF_SLOTS; ++i) { ms[i] = e->prof_ms[i]; counts[i] = e->prof_count[i]; }
    return SVO_OK;
}
int svo_timer_begin(svo_engine* e) {
    if (!e) return SVO_ERR_INVALID;
    ECUDA(e, cudaSetDevice(e->device));
    ECUDA(e, cudaStreamSynchronize(e->st));
    ECUDA(e, cudaEventRecord(e->timer_ev[0], e->st));
    return SVO_OK;
}
int svo_timer_end(svo_engine* e, float* ms) {
    if (!e || !ms) return SVO_ERR_INVALID;
    ECUDA(e, cudaSetDevice(e->device));
    ECUDA(e, cudaEventRecord(e->timer_ev[1], e->st));
    ECUDA(e, cudaEventSynchronize(e->timer_ev[1]));
    ECUDA(e, cudaEventElapsedTime(ms, e->timer_ev[0], e->timer_ev[1]));
    return SVO_OK;
}

int svo_step(svo_engine* e, int32_t n, const int32_t* seqs, const uint8_t* const* left,
             const uint8_t* const* right, int32_t pitch, svo_float2* corners, int32_t* n_corners,
             svo_track_stats* stats, svo_float2* tracked, uint64_t* tracked_ids) {
    return step_impl(e, n, seqs, left, right, pitch, false, corners, n_corners, stats, tracked,
                     tracked_ids);
}

int svo_prefetch(svo_engine* e, int32_t n, const int32_t* seqs, const uint8_t* const* left,
                 const uint8_t* const* right, int32_t pitch) {
    if (!e) return SVO_ERR_INVALID;
    if (n <= 0 || n > e->n_seq || !seqs || !left) return fail(e, SVO_ERR_INVALID, "svo_prefetch: bad arguments");
    if (pitch < e->p.width) return fail(e, SVO_ERR_INVALID, "svo_prefetch: pitch < width");
    if (e->pf_count >= 2) return fail(e, SVO_ERR_STATE, "svo_prefetch: two prefetches already pending");
    ECUDA(e, cudaSetDevice(e->device));
    // the set the other pending entry (if any) does not use; with none pending, not the set of
    // the step that may still be running its ingest kernel
    int set = e->pf_count ? 1 - e->pf[e->pf_head].set : 1 - e->cur_set;
    svo_engine::Prefetch& p = e->pf[e->pf_head ^ (e->pf_count ? 1 : 0)];
    p.seqs.assign(seqs, seqs + n);
    p.left.assign(left, left + n);
    p.right.assign((size_t)n, nullptr);
    for (int i = 0; i < n; ++i) {
        if (seqs[i] < 0 || seqs[i] >= e->n_seq) return fail(e, SVO_ERR_INVALID, "svo_prefetch: sequence index out of range");
        if (right) p.right[i] = right[i];
    }
    ECUDA(e, stage_batch(e, e->st_copy, set, n, seqs, p.left.data(), p.right.data(), pitch));
    ECUDA(e, cudaEventRecord(p.done, e->st_copy));
    p.valid = true;
    p.skipped = 0;
    p.set = set;
    p.n = n;
    p.pitch = pitch;
    e->pf_count += 1;
    return SVO_OK;
}

int svo_step_device(svo_engine* e, int32_t n, const int32_t* seqs, const uint8_t* const* d_left,
                    const uint8_t* const* d_right, int32_t pitch, svo_float2* corners,
                    int32_t* n_corners, svo_track_stats* stats, svo_float2* tracked,
                    uint64_t* tracked_ids) {
    return step_impl(e, n, seqs, d_left, d_right, pitch, true, corners, n_corners, stats, tracked,
                     tracked_ids);
}

int svo_good_features_to_track(svo_engine* e, int32_t seq, const uint8_t* img, int32_t pitch,
                               svo_float2* out_xy, int32_t* out_n) {
    int rc = check_seq(e, seq);
    if (rc) return rc;
    if (!img || !out_xy || !out_n || pitch < e->p.width) return fail(e, SVO_ERR_INVALID, "svo_good_features_to_track: bad arguments");
    ECUDA(e, cudaSetDevice(e->device));
    begin_frame(e, seq);
    SvoSeqCall* hc = next_call_slot(e);
    fill_call(e, hc[0], seq);
    ECUDA(e, upload_left(e, hc[0], img, pitch, false));
    ECUDA(e, cudaEventRecord(e->ev[0], e->st));
    ECUDA(e, cudaEventRecord(e->ev[1], e->st));
    ECUDA(e, push_and_ingest(e, 1));
    ECUDA(e, launch_detect(e, 1, true));
    ECUDA(e, cudaEventRecord(e->ev[2], e->st));
    int hdr[2] = {0, 0};
    ECUDA(e, cudaMemcpyAsync(&hdr[0], e->d_ncorners + seq, sizeof(int), cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaMemcpyAsync(&hdr[1], e->d_cand_count + seq, sizeof(int), cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaMemcpyAsync(e->h_out, e->d_corners + (size_t)seq * e->p.max_corners,
                             (size_t)e->p.max_corners * sizeof(float2), cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaEventRecord(e->ev[3], e->st));
    ECUDA(e, wait_stream(e));
    finish_timing(e);
    if (hdr[1] > e->max_cand) {
        // plateau image: repeat on the spill lists (room for every pixel)
        rc = detect_spill(e, seq);
        if (rc) return rc;
        ECUDA(e, cudaMemcpyAsync(&hdr[0], e->d_ncorners + seq, sizeof(int), cudaMemcpyDeviceToHost, e->st));
        ECUDA(e, cudaMemcpyAsync(&hdr[1], e->d_cand_count + seq, sizeof(int), cudaMemcpyDeviceToHost, e->st));
        ECUDA(e, cudaMemcpyAsync(e->h_out, e->d_corners + (size_t)seq * e->p.max_corners,
                                 (size_t)e->p.max_corners * sizeof(float2), cudaMemcpyDeviceToHost, e->st));
        ECUDA(e, wait_stream(e));
        if (hdr[1] > e->max_cand_big) return fail(e, SVO_ERR_CAPACITY, "detector spill lists overflowed (internal error)");
    }
    *out_n = hdr[0];
    memcpy(out_xy, e->h_out, (size_t)hdr[0] * sizeof(float2));
    return SVO_OK;
}

static int keyframe_impl(svo_engine* e, int n, const int32_t* seqs, const svo_float2* inliers,
                         const int32_t* n_inliers, const float* poses, float focal, float cx,
                         float cy, float baseline, bool dedupe, const svo_float2* explicit_pts,
                         const int32_t* n_explicit, svo_float2* new_feats, int32_t* n_new,
                         svo_float2* valid_2d, svo_float3* pts_3d, int32_t* valid_idx,
                         int16_t* disp16, int32_t* n_valid,
                         const uint64_t* first_new_id = nullptr) {
    // first_new_id != nullptr: every tracked feature of the sequence is an inlier - the list is
    // the tracker's device-resident feature set, and the tracker is re-initialised on the
    // device at the end (nothing but the descriptors travels to the GPU)
    const bool tracked = first_new_id != nullptr;
    const int mf = e->p.max_features, cap = e->pts_cap, mcn = e->p.max_corners;
    ECUDA(e, cudaSetDevice(e->device));
    SvoSeqCall* hc = next_call_slot(e);
    float2* h_pts = reinterpret_cast<float2*>(e->h_in);
    int max_pts = 0;
    for (int i = 0; i < n; ++i) {
        int s = seqs[i];
        if (s < 0 || s >= e->n_seq) return fail(e, SVO_ERR_INVALID, "sequence index out of range");
        if (!e->seq[s].have_frame || !e->seq[s].have_right)
            return fail(e, SVO_ERR_STATE, "stereo stage needs an uploaded stereo pair");
        fill_call(e, hc[i], s);
        if (poses) memcpy(hc[i].pose, poses + (size_t)i * 16, 16 * sizeof(float));
        if (tracked) {
            hc[i].n_aux = e->seq[s].tracker_init ? e->seq[s].n_feat : 0;
            hc[i].n_pts = 0;
            hc[i].first_id = first_new_id[i];
            max_pts = mcn;
        } else if (dedupe) {
            int ni = n_inliers ? n_inliers[i] : 0;
            if (ni < 0 || ni > mf) return fail(e, SVO_ERR_INVALID, "n_inliers out of range");
            hc[i].n_aux = ni;
            hc[i].n_pts = 0;
            if (ni > 0) memcpy(h_pts + (size_t)i * mf, inliers + (size_t)i * mf, (size_t)ni * sizeof(float2));
            max_pts = mcn;
        } else {
            int np = n_explicit[i];
            if (np < 0 || np > cap) return fail(e, SVO_ERR_CAPACITY, "too many features for the stereo stage");
            hc[i].n_pts = np;
            if (np > max_pts) max_pts = np;
            if (np > 0) memcpy(h_pts + (size_t)i * cap, explicit_pts + (size_t)i * cap, (size_t)np * sizeof(float2));
        }
    }
    ECUDA(e, cudaEventRecord(e->ev[0], e->st));
    SvoSeqCall* dc = dev_call_slot(e);
    if (tracked) {
        ECUDA(e, push_calls(e, n));
    } else {
        PullList pl(e);
        pl.add(dc, hc, (size_t)n * sizeof(SvoSeqCall));
        for (int i = 0; i < n; ++i) {
            int s = seqs[i];
            if (dedupe)
                pl.add(e->d_inl + (size_t)s * mf, h_pts + (size_t)i * mf, (size_t)hc[i].n_aux * sizeof(float2));
            else
                pl.add(e->d_pts + (size_t)s * cap, h_pts + (size_t)i * cap, (size_t)hc[i].n_pts * sizeof(float2));
        }
        ECUDA(e, pl.launch());
    }
    ECUDA(e, cudaEventRecord(e->ev[1], e->st));
    if (dedupe) {
        ProfScope ps(e, P_DEDUPE);
        ECUDA(e, svo_launch_dedupe(e->d_corners, e->d_ncorners, mcn, tracked ? e->d_feat : e->d_inl,
                                   tracked ? 2 * mf : mf, tracked ? mf : 0, e->p.min_distance,
                                   e->d_pts, e->d_keep_idx, e->d_nnew, cap, dc, e->d_ticket, n, e->st));
        e->launches += 1;
    }
    SvoStereoParams sp = {e->p.num_disparities, e->p.sad_window, e->p.prefilter_cap,
                          e->p.texture_threshold, e->p.uniqueness_ratio};
    if (max_pts > 0) {
        ProfScope ps(e, P_STEREO);
        ECUDA(e, svo_launch_stereo_bm_sparse(e->d_pyr, e->lay, e->seq_stride, e->d_right,
                                             e->right_stride, e->d_pts, cap, e->d_disp16, sp,
                                             max_pts, dc, n, e->st));
        e->launches += 1;
    }
    {
        ProfScope ps(e, P_REPROJ);
        ECUDA(e, svo_launch_reproject(e->d_pts, e->d_disp16, cap, focal, cx, cy, baseline,
                                      e->d_valid2d, e->d_pts3d, e->d_valid_idx, e->d_nvalid, dc, n,
                                      e->st));
    }
    if (tracked) {
        tracker_append_kernel<<<n, 128, 0, e->st>>>(e->d_feat, e->d_ids, e->d_initial, mf,
                                                    e->d_valid2d, cap, e->d_nvalid, dc);
        e->launches += 1;
    }
    pack_key_kernel<<<n, 128, 0, e->st>>>(e->d_out, e->key_block, cap, e->d_pts,
                                          dedupe ? e->d_nnew : nullptr, e->d_valid2d, e->d_pts3d,
                                          e->d_valid_idx, e->d_disp16, e->d_nvalid, dc);
    e->launches += 2;
    ECUDA(e, cudaGetLastError());
    ECUDA(e, cudaEventRecord(e->ev[2], e->st));
    ECUDA(e, cudaMemcpyAsync(e->h_out, e->d_out, (size_t)n * e->key_block, cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaEventRecord(e->ev[3], e->st));
    ECUDA(e, wait_stream(e));
    finish_timing(e);
    for (int i = 0; i < n; ++i) {
        const unsigned char* o = e->h_out + (size_t)i * e->key_block;
        const KeyHeader* h = reinterpret_cast<const KeyHeader*>(o);
        const float2* o_new = reinterpret_cast<const float2*>(o + sizeof(KeyHeader));
        const float2* o_v2 = o_new + cap;
        const float* o_3d = reinterpret_cast<const float*>(o_v2 + cap);
        const int* o_vi = reinterpret_cast<const int*>(o_3d + 3 * (size_t)cap);
        const short* o_d = reinterpret_cast<const short*>(o_vi + cap);
        const size_t ob = (size_t)i * (dedupe ? mcn : cap);
        if (n_new) n_new[i] = h->n_new;
        if (n_valid) n_valid[i] = h->n_valid;
        if (new_feats) memcpy(new_feats + ob, o_new, (size_t)h->n_new * sizeof(float2));
        if (disp16) memcpy(disp16 + ob, o_d, (size_t)h->n_new * sizeof(short));
        if (valid_2d) memcpy(valid_2d + ob, o_v2, (size_t)h->n_valid * sizeof(float2));
        if (valid_idx) memcpy(valid_idx + ob, o_vi, (size_t)h->n_valid * sizeof(int));
        if (pts_3d) memcpy(pts_3d + ob, o_3d, (size_t)h->n_valid * 3 * sizeof(float));
        if (tracked) {                      // the host mirror of what tracker_append_kernel did
            SeqState& st = e->seq[seqs[i]];
            const int nt = hc[i].n_aux;
            st.n_feat = nt + std::min((int)h->n_valid, mf - nt);
            st.n_initial = st.n_feat;
            st.tracker_init = true;
            st.last_slot = st.cur_slot;    // last_image = image.clone()  (feature_tracker.cpp:14)
        }
    }
    return SVO_OK;
}

int svo_step_keyframe(svo_engine* e, int32_t n, const int32_t* seqs, const uint8_t* const* left,
                      const uint8_t* const* right, int32_t pitch, int32_t on_device,
                      const float* poses, const uint64_t* first_new_id, float parallax_thresh,
                      float focal, float cx, float cy, float baseline, svo_float2* corners,
                      int32_t* n_corners, svo_track_stats* stats, svo_float2* tracked,
                      uint64_t* tracked_ids, int32_t* decision, svo_float2* new_feats,
                      int32_t* n_new, svo_float2* valid_2d, svo_float3* pts_3d,
                      int32_t* valid_idx, int16_t* disp16, int32_t* n_valid) {
    if (!e) return SVO_ERR_INVALID;
    if (!poses || !first_new_id || !decision || !n_new || !n_valid)
        return fail(e, SVO_ERR_INVALID, "svo_step_keyframe: bad arguments");
    const StepKey key = {poses, first_new_id, parallax_thresh, focal, cx, cy, baseline, decision,
                         new_feats, n_new, valid_2d, pts_3d, valid_idx, disp16, n_valid};
    return step_impl(e, n, seqs, left, right, pitch, on_device != 0, corners, n_corners, stats,
                     tracked, tracked_ids, &key);
}

int svo_keyframe_tracked(svo_engine* e, int32_t n, const int32_t* seqs, const float* poses,
                         const uint64_t* first_new_id, float focal, float cx, float cy,
                         float baseline, svo_float2* new_feats, int32_t* n_new,
                         svo_float2* valid_2d, svo_float3* pts_3d, int32_t* valid_idx,
                         int16_t* disp16, int32_t* n_valid) {
    if (!e) return SVO_ERR_INVALID;
    if (n <= 0 || n > e->n_seq || !seqs || !poses || !first_new_id)
        return fail(e, SVO_ERR_INVALID, "svo_keyframe_tracked: bad arguments");
    return keyframe_impl(e, n, seqs, nullptr, nullptr, poses, focal, cx, cy, baseline, true,
                         nullptr, nullptr, new_feats, n_new, valid_2d, pts_3d, valid_idx, disp16,
                         n_valid, first_new_id);
}

int svo_keyframe(svo_engine* e, int32_t n, const int32_t* seqs, const svo_float2* inliers,
                 const int32_t* n_inliers, const float* poses, float focal, float cx, float cy,
                 float baseline, svo_float2* new_feats, int32_t* n_new, svo_float2* valid_2d,
                 svo_float3* pts_3d, int32_t* valid_idx, int16_t* disp16, int32_t* n_valid) {
    if (!e) return SVO_ERR_INVALID;
    if (n <= 0 || n > e->n_seq || !seqs || !poses) return fail(e, SVO_ERR_INVALID, "svo_keyframe: bad arguments");
    return keyframe_impl(e, n, seqs, inliers, n_inliers, poses, focal, cx, cy, baseline, true,
                         nullptr, nullptr, new_feats, n_new, valid_2d, pts_3d, valid_idx, disp16,
                         n_valid);
}

int svo_triangulate_stereo(svo_engine* e, int32_t seq, const uint8_t* left, const uint8_t* right,
                           int32_t pitch, const svo_float2* feats, int32_t n, const float pose[16],
                           float focal, float cx, float cy, float baseline, svo_float3* out_3d,
                           svo_float2* out_valid_2d, int32_t* out_valid_idx, int16_t* out_disp16,
                           int32_t* out_n_valid) {
    int rc = check_seq(e, seq);
    if (rc) return rc;
    if (!left || !right || (!feats && n > 0) || !pose || n < 0 || pitch < e->p.width)
        return fail(e, SVO_ERR_INVALID, "svo_triangulate_stereo: bad arguments");
    if (n > e->pts_cap) return fail(e, SVO_ERR_CAPACITY, "svo_triangulate_stereo: more features than max(max_corners, max_features)");
    ECUDA(e, cudaSetDevice(e->device));
    begin_frame(e, seq);
    {
        SvoSeqCall* hc = next_call_slot(e);
        fill_call(e, hc[0], seq);
        ECUDA(e, upload_left(e, hc[0], left, pitch, false));
        ECUDA(e, upload_right(e, hc[0], right, pitch, false));
        ECUDA(e, push_and_ingest(e, 1));
    }
    // single-sequence call: element 0 of the [n][cap] arrays
    int32_t s = seq, nn = n;
    return keyframe_impl(e, 1, &s, nullptr, nullptr, pose, focal, cx, cy, baseline, false, feats,
                         &nn, nullptr, nullptr, out_valid_2d, out_3d, out_valid_idx, out_disp16,
                         out_n_valid);
}

// Dense StereoBM over the stereo pairs that are on the device (left = current pyramid level 0).
static int dense_impl(svo_engine* e, int n, const int32_t* seqs, int16_t* const* out, int32_t out_pitch) {
    const int W = e->p.width, H = e->p.height;
    SvoStereoParams sp = {e->p.num_disparities, e->p.sad_window, e->p.prefilter_cap,
                          e->p.texture_threshold, e->p.uniqueness_ratio};
    if (!svo_stereo_dense_supported(sp))
        return fail(e, SVO_ERR_INVALID, "svo_stereo_bm_dense: built for StereoBM(48, 21), prefilter cap 1..63");
    if (out && out_pitch < W) return fail(e, SVO_ERR_INVALID, "svo_stereo_bm_dense: out_pitch < width");
    const size_t map = (size_t)W * H;
    if (!e->d_dense) {
        if (dalloc(&e->d_dense, (size_t)e->n_seq * map) != cudaSuccess) {
            cudaGetLastError();
            e->d_dense = nullptr;
            return fail(e, SVO_ERR_NOMEM, "svo_stereo_bm_dense: no device memory for the disparity maps");
        }
    }
    SvoSeqCall* hc = next_call_slot(e);
    for (int i = 0; i < n; ++i) {
        const int s = seqs[i];
        if (s < 0 || s >= e->n_seq) return fail(e, SVO_ERR_INVALID, "sequence index out of range");
        if (!e->seq[s].have_frame || !e->seq[s].have_right)
            return fail(e, SVO_ERR_STATE, "svo_stereo_bm_dense: needs an uploaded stereo pair");
        fill_call(e, hc[i], s);
    }
    ECUDA(e, cudaEventRecord(e->ev[0], e->st));
    SvoSeqCall* dc = dev_call_slot(e);
    ECUDA(e, push_calls(e, n));
    ECUDA(e, cudaEventRecord(e->ev[1], e->st));
    {
        ProfScope ps(e, P_DENSE);
        ECUDA(e, svo_launch_stereo_bm_dense(e->d_pyr, e->lay, e->seq_stride, e->d_right,
                                            e->right_stride, e->d_dense, W, map, sp, dc, n, e->st));
        e->launches += 2;
    }
    ECUDA(e, cudaEventRecord(e->ev[2], e->st));
    for (int i = 0; out && i < n; ++i)
        if (out[i])
            ECUDA(e, cudaMemcpy2DAsync(out[i], (size_t)out_pitch * sizeof(int16_t),
                                       e->d_dense + (size_t)seqs[i] * map, (size_t)W * sizeof(int16_t),
                                       (size_t)W * sizeof(int16_t), H, cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaEventRecord(e->ev[3], e->st));
    ECUDA(e, wait_stream(e));
    finish_timing(e);
    return SVO_OK;
}

int svo_stereo_bm_dense(svo_engine* e, int32_t seq, const uint8_t* left, const uint8_t* right,
                        int32_t pitch, int16_t* out_disp16, int32_t out_pitch) {
    int rc = check_seq(e, seq);
    if (rc) return rc;
    if (!left || !right || !out_disp16 || pitch < e->p.width)
        return fail(e, SVO_ERR_INVALID, "svo_stereo_bm_dense: bad arguments");
    ECUDA(e, cudaSetDevice(e->device));
    begin_frame(e, seq);
    {
        SvoSeqCall* hc = next_call_slot(e);
        fill_call(e, hc[0], seq);
        ECUDA(e, upload_left(e, hc[0], left, pitch, false));
        ECUDA(e, upload_right(e, hc[0], right, pitch, false));
        ECUDA(e, push_and_ingest(e, 1));
    }
    int32_t s = seq;
    int16_t* o = out_disp16;
    return dense_impl(e, 1, &s, &o, out_pitch);
}

int svo_stereo_bm_dense_resident(svo_engine* e, int32_t n, const int32_t* seqs,
                                 int16_t* const* out_disp16, int32_t out_pitch) {
    if (!e) return SVO_ERR_INVALID;
    if (n <= 0 || n > e->n_seq || !seqs) return fail(e, SVO_ERR_INVALID, "svo_stereo_bm_dense_resident: bad arguments");
    ECUDA(e, cudaSetDevice(e->device));
    return dense_impl(e, n, seqs, out_disp16, out_pitch);
}

int svo_pnp_ransac(svo_engine* e, int32_t n, const svo_float3* world, const svo_float2* image,
                   const int32_t* counts, float focal, float cx, float cy, int32_t iterations,
                   float reproj_error, double confidence, const double* guess,
                   svo_pnp_result* results, int32_t* inliers) {
    static_assert(sizeof(svo_pnp_result) == sizeof(SvoPnpResult), "result layouts must match");
    if (!e) return SVO_ERR_INVALID;
    if (n <= 0 || n > e->n_seq || !world || !image || !counts || !guess || !results || !inliers)
        return fail(e, SVO_ERR_INVALID, "svo_pnp_ransac: bad arguments");
    if (iterations <= 0 || iterations > 128 || !(reproj_error > 0.f) || !(confidence > 0.0))
        return fail(e, SVO_ERR_INVALID, "svo_pnp_ransac: iterations 1..128, positive threshold and confidence");
    ECUDA(e, cudaSetDevice(e->device));
    const int mf = e->p.max_features;
    const size_t ns = (size_t)e->n_seq;
    if (!e->d_pnp_world) {
        if (dalloc(&e->d_pnp_world, ns * mf * 3) != cudaSuccess || dalloc(&e->d_pnp_img, ns * mf) != cudaSuccess ||
            dalloc(&e->d_pnp_counts, ns) != cudaSuccess || dalloc(&e->d_pnp_guess, ns * 6) != cudaSuccess ||
            dalloc(&e->d_pnp_res, ns) != cudaSuccess || dalloc(&e->d_pnp_inl, ns * mf) != cudaSuccess) {
            cudaGetLastError();
            return fail(e, SVO_ERR_NOMEM, "svo_pnp_ransac: no device memory");
        }
    }
    for (int i = 0; i < n; ++i)
        if (counts[i] < 0 || counts[i] > mf) return fail(e, SVO_ERR_CAPACITY, "svo_pnp_ransac: more points than max_features");
    ECUDA(e, cudaEventRecord(e->ev[0], e->st));
    for (int i = 0; i < n; ++i) {
        if (counts[i] == 0) continue;
        ECUDA(e, cudaMemcpyAsync(e->d_pnp_world + (size_t)i * mf * 3, world + (size_t)i * mf,
                                 (size_t)counts[i] * 3 * sizeof(float), cudaMemcpyHostToDevice, e->st));
        ECUDA(e, cudaMemcpyAsync(e->d_pnp_img + (size_t)i * mf, image + (size_t)i * mf,
                                 (size_t)counts[i] * sizeof(float2), cudaMemcpyHostToDevice, e->st));
    }
    ECUDA(e, cudaMemcpyAsync(e->d_pnp_counts, counts, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, e->st));
    ECUDA(e, cudaMemcpyAsync(e->d_pnp_guess, guess, (size_t)n * 6 * sizeof(double), cudaMemcpyHostToDevice, e->st));
    ECUDA(e, cudaEventRecord(e->ev[1], e->st));
    {
        ProfScope ps(e, P_PNP);
        ECUDA(e, svo_launch_pnp_ransac(e->d_pnp_world, e->d_pnp_img, mf, e->d_pnp_counts,
                                       e->d_pnp_guess, focal, cx, cy, iterations, reproj_error,
                                       confidence, e->pnp_calls++, e->d_pnp_res, e->d_pnp_inl, n, e->st));
        e->launches += 1;
    }
    ECUDA(e, cudaEventRecord(e->ev[2], e->st));
    ECUDA(e, cudaMemcpyAsync(results, e->d_pnp_res, (size_t)n * sizeof(SvoPnpResult), cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaMemcpyAsync(inliers, e->d_pnp_inl, (size_t)n * mf * sizeof(int), cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaEventRecord(e->ev[3], e->st));
    ECUDA(e, wait_stream(e));
    finish_timing(e);
    return SVO_OK;
}

int svo_reprojection_evaluate(svo_engine* e, int32_t n_obs, const int32_t* pose_idx,
                              const int32_t* point_idx, const double* obs, int32_t n_poses,
                              const double* poses, int32_t n_points, const double* points,
                              double focal, double cx, double cy, double* residuals,
                              double* jac_pose, double* jac_point) {
    if (!e) return SVO_ERR_INVALID;
    if (n_obs < 0 || n_poses <= 0 || n_points <= 0 || !pose_idx || !point_idx || !obs || !poses || !points)
        return fail(e, SVO_ERR_INVALID, "svo_reprojection_evaluate: bad arguments");
    for (int i = 0; i < n_obs; ++i)
        if (pose_idx[i] < 0 || pose_idx[i] >= n_poses || point_idx[i] < 0 || point_idx[i] >= n_points)
            return fail(e, SVO_ERR_INVALID, "svo_reprojection_evaluate: index out of range");
    if (n_obs == 0) return SVO_OK;
    ECUDA(e, cudaSetDevice(e->device));
    // one scratch allocation per call shape (a window is rebuilt once per keyframe, not per frame)
    const size_t b_poses = (size_t)n_poses * 7 * 8, b_pts = (size_t)n_points * 3 * 8,
                 b_idx = (size_t)n_obs * 4, b_obs = (size_t)n_obs * 2 * 8;
    const size_t b_out = (size_t)n_obs * (2 + 14 + 6) * 8;
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };       // slices start 256-B aligned
    const size_t need = up(b_poses) + up(b_pts) + 2 * up(b_idx) + up(b_obs) + b_out;
    if (need > e->ba_bytes) {
        if (e->d_ba) cudaFree(e->d_ba);
        e->d_ba = nullptr;
        e->ba_bytes = 0;
        if (cudaMalloc(reinterpret_cast<void**>(&e->d_ba), need) != cudaSuccess) {
            cudaGetLastError();
            return fail(e, SVO_ERR_NOMEM, "svo_reprojection_evaluate: no device memory");
        }
        e->ba_bytes = need;
    }
    unsigned char* base = e->d_ba;
    double* d_poses = reinterpret_cast<double*>(base); base += up(b_poses);
    double* d_pts = reinterpret_cast<double*>(base); base += up(b_pts);
    double* d_obs = reinterpret_cast<double*>(base); base += up(b_obs);
    int* d_pi = reinterpret_cast<int*>(base); base += up(b_idx);
    int* d_xi = reinterpret_cast<int*>(base); base += up(b_idx);
    double* d_r = reinterpret_cast<double*>(base);
    double* d_jp = d_r + (size_t)n_obs * 2;
    double* d_jx = d_jp + (size_t)n_obs * 14;
    ECUDA(e, cudaEventRecord(e->ev[0], e->st));
    ECUDA(e, cudaMemcpyAsync(d_poses, poses, b_poses, cudaMemcpyHostToDevice, e->st));
    ECUDA(e, cudaMemcpyAsync(d_pts, points, b_pts, cudaMemcpyHostToDevice, e->st));
    ECUDA(e, cudaMemcpyAsync(d_obs, obs, b_obs, cudaMemcpyHostToDevice, e->st));
    ECUDA(e, cudaMemcpyAsync(d_pi, pose_idx, b_idx, cudaMemcpyHostToDevice, e->st));
    ECUDA(e, cudaMemcpyAsync(d_xi, point_idx, b_idx, cudaMemcpyHostToDevice, e->st));
    ECUDA(e, cudaEventRecord(e->ev[1], e->st));
    ECUDA(e, svo_launch_reprojection_factor(d_poses, d_pts, d_pi, d_xi, d_obs, n_obs, focal, cx, cy,
                                            d_r, d_jp, d_jx, e->st));
    e->launches += 1;
    ECUDA(e, cudaEventRecord(e->ev[2], e->st));
    if (residuals) ECUDA(e, cudaMemcpyAsync(residuals, d_r, (size_t)n_obs * 2 * 8, cudaMemcpyDeviceToHost, e->st));
    if (jac_pose) ECUDA(e, cudaMemcpyAsync(jac_pose, d_jp, (size_t)n_obs * 14 * 8, cudaMemcpyDeviceToHost, e->st));
    if (jac_point) ECUDA(e, cudaMemcpyAsync(jac_point, d_jx, (size_t)n_obs * 6 * 8, cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaEventRecord(e->ev[3], e->st));
    ECUDA(e, wait_stream(e));
    finish_timing(e);
    return SVO_OK;
}

int svo_draw_track(svo_engine* e, const uint8_t* image, int32_t pitch, const svo_float2* from,
                   const svo_float2* to, int32_t n, uint8_t* out_rgb, int32_t out_pitch) {
    if (!e) return SVO_ERR_INVALID;
    const int W = e->p.width, H = e->p.height;
    if (!image || !out_rgb || pitch < W || out_pitch < 3 * W || n < 0 || (n > 0 && (!from || !to)))
        return fail(e, SVO_ERR_INVALID, "svo_draw_track: bad arguments");
    if (n > e->p.max_features) return fail(e, SVO_ERR_CAPACITY, "svo_draw_track: more features than max_features");
    ECUDA(e, cudaSetDevice(e->device));
    const size_t gray_b = (size_t)W * H, rgb_b = 3 * gray_b, pts_b = (size_t)e->p.max_features * sizeof(float2);
    if (!e->d_draw) {
        if (cudaMalloc(reinterpret_cast<void**>(&e->d_draw), gray_b + rgb_b + 2 * pts_b + 1024) != cudaSuccess) {
            cudaGetLastError();
            return fail(e, SVO_ERR_NOMEM, "svo_draw_track: no device memory");
        }
    }
    uint8_t* d_rgb = e->d_draw;
    uint8_t* d_gray = d_rgb + ((rgb_b + 255) & ~(size_t)255);
    float2* d_from = reinterpret_cast<float2*>(d_gray + ((gray_b + 255) & ~(size_t)255));
    float2* d_to = d_from + e->p.max_features;
    ECUDA(e, cudaMemcpy2DAsync(d_gray, W, image, pitch, W, H, cudaMemcpyHostToDevice, e->st));
    if (n > 0) {
        ECUDA(e, cudaMemcpyAsync(d_from, from, (size_t)n * sizeof(float2), cudaMemcpyHostToDevice, e->st));
        ECUDA(e, cudaMemcpyAsync(d_to, to, (size_t)n * sizeof(float2), cudaMemcpyHostToDevice, e->st));
    }
    ECUDA(e, svo_launch_draw_track(d_gray, W, d_from, d_to, n, d_rgb, 3 * W, W, H, 4, e->st));
    e->launches += n > 0 ? 2 : 1;
    ECUDA(e, cudaMemcpy2DAsync(out_rgb, out_pitch, d_rgb, 3 * (size_t)W, 3 * (size_t)W, H,
                               cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaStreamSynchronize(e->st));
    return SVO_OK;
}

int svo_device_can_read(const void* p) {
    if (!p) return 0;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) return 1;
    // pinned host memory is mapped into the device address space under unified addressing;
    // kernels can only use it through the same pointer value
    return (a.type == cudaMemoryTypeHost && a.devicePointer == p) ? 1 : 0;
}

int svo_set_right(svo_engine* e, int32_t n, const int32_t* seqs, const uint8_t* const* right,
                  int32_t pitch, int32_t on_device) {
    if (!e) return SVO_ERR_INVALID;
    if (n <= 0 || n > e->n_seq || !seqs || !right) return fail(e, SVO_ERR_INVALID, "svo_set_right: bad arguments");
    if (pitch < e->p.width) return fail(e, SVO_ERR_INVALID, "svo_set_right: pitch < width");
    ECUDA(e, cudaSetDevice(e->device));
    if (!on_device) {
        // land in a staging set no announced (prefetched) frames occupy
        if (e->pf_count == 2) prefetch_discard(e);
        if (e->pf_count == 1) e->cur_set = 1 - e->pf[e->pf_head].set;
    }
    SvoSeqCall* hc = next_call_slot(e);
    for (int i = 0; i < n; ++i) {
        const int s = seqs[i];
        if (s < 0 || s >= e->n_seq) return fail(e, SVO_ERR_INVALID, "svo_set_right: sequence index out of range");
        if (!right[i]) return fail(e, SVO_ERR_INVALID, "svo_set_right: null image");
        if (!e->seq[s].have_frame) return fail(e, SVO_ERR_STATE, "svo_set_right: no current left frame");
        fill_call(e, hc[i], s);
        ECUDA(e, upload_right(e, hc[i], right[i], pitch, on_device != 0));
    }
    ECUDA(e, push_and_ingest(e, n));      // asynchronous: ordered before the next call's kernels
    return SVO_OK;
}

int svo_tracker_set(svo_engine* e, int32_t n, const int32_t* seqs, const svo_float2* feats,
                    const uint64_t* ids, const int32_t* counts) {
    if (!e) return SVO_ERR_INVALID;
    if (n <= 0 || n > e->n_seq || !seqs || !counts) return fail(e, SVO_ERR_INVALID, "svo_tracker_set: bad arguments");
    ECUDA(e, cudaSetDevice(e->device));
    const int mf = e->p.max_features;
    // The call does not wait for the device: the feature lists are staged in a buffer of their
    // own and every later call on this engine is ordered behind the copy on the engine stream.
    // Only a second svo_tracker_set has to wait until the first one's staging has been read.
    ECUDA(e, cudaEventSynchronize(e->ev_trk));
    float2* hf = reinterpret_cast<float2*>(e->h_trk);
    unsigned long long* hi = reinterpret_cast<unsigned long long*>(hf + (size_t)e->n_seq * mf);
    for (int i = 0; i < n; ++i) {
        int s = seqs[i];
        if (s < 0 || s >= e->n_seq) return fail(e, SVO_ERR_INVALID, "sequence index out of range");
        if (!e->seq[s].have_frame) return fail(e, SVO_ERR_STATE, "svo_tracker_set: no current frame");
        int c = counts[i];
        if (c < 0 || c > mf) return fail(e, SVO_ERR_CAPACITY, "svo_tracker_set: more features than max_features");
        if (c > 0 && (!feats || !ids)) return fail(e, SVO_ERR_INVALID, "svo_tracker_set: null arrays");
        if (c > 0) {
            memcpy(hf + (size_t)i * mf, feats + (size_t)i * mf, (size_t)c * sizeof(float2));
            memcpy(hi + (size_t)i * mf, ids + (size_t)i * mf, (size_t)c * sizeof(uint64_t));
        }
    }
    PullList pl(e);
    for (int i = 0; i < n; ++i) {
        SeqState& s = e->seq[seqs[i]];
        int c = counts[i];
        const size_t fb = ((size_t)seqs[i] * 2 + s.feat_buf) * mf;
        pl.add(e->d_feat + fb, hf + (size_t)i * mf, (size_t)c * sizeof(float2));
        pl.add(e->d_initial + fb, hf + (size_t)i * mf, (size_t)c * sizeof(float2));
        pl.add(e->d_ids + fb, hi + (size_t)i * mf, (size_t)c * sizeof(uint64_t));
        s.n_feat = c;
        // initial_features is an unordered_map keyed by id: duplicates collapse (feature_tracker.cpp:10-12)
        int uniq = 0;
        if (c > 0) {
            std::vector<uint64_t> v(ids + (size_t)i * mf, ids + (size_t)i * mf + c);
            std::sort(v.begin(), v.end());
            uniq = (int)(std::unique(v.begin(), v.end()) - v.begin());
        }
        s.n_initial = uniq;
        s.tracker_init = true;
        s.last_slot = s.cur_slot;      // last_image = image.clone()  (feature_tracker.cpp:14)
    }
    ECUDA(e, pl.launch());
    if (e->lk_handover && e->lk_rekey) {
        // point the hand-over index map at the entries of the features the host kept
        // (lk_track.cu: lk_rekey_kernel), 32 sequences per launch, behind the copy of the lists;
        // launch_track waits for it before the next forward pass
        const SvoLkCache lkc = {e->d_lk_data, e->d_lk_hdr, e->d_img_serial, e->d_kept_idx, nullptr};
        ECUDA(e, cudaEventRecord(e->ev_feat, e->st));
        ECUDA(e, cudaStreamWaitEvent(e->st_rekey, e->ev_feat, 0));
        for (int i0 = 0; i0 < n; i0 += 32) {
            SvoLkRekey rk;
            rk.n = std::min(32, n - i0);
            for (int i = 0; i < rk.n; ++i) {
                const SeqState& s = e->seq[seqs[i0 + i]];
                rk.seq[i] = seqs[i0 + i];
                rk.slot[i] = s.cur_slot;
                rk.feat_buf[i] = s.feat_buf;
                rk.count[i] = counts[i0 + i];
            }
            ECUDA(e, svo_launch_lk_rekey(e->d_feat, mf, rk, lkc, e->d_kept_idx, e->st_rekey));
            e->launches += 1;
        }
        ECUDA(e, cudaEventRecord(e->ev_rekey, e->st_rekey));
        e->rekey_pending = true;
    }
    ECUDA(e, cudaEventRecord(e->ev_trk, e->st));
    return SVO_OK;
}

int svo_tracker_init(svo_engine* e, int32_t seq, const uint8_t* img, int32_t pitch,
                     const svo_float2* feats, const uint64_t* ids, int32_t n) {
    int rc = check_seq(e, seq);
    if (rc) return rc;
    if (!img || pitch < e->p.width || n < 0) return fail(e, SVO_ERR_INVALID, "svo_tracker_init: bad arguments");
    if (n > e->p.max_features) return fail(e, SVO_ERR_CAPACITY, "svo_tracker_init: more features than max_features");
    ECUDA(e, cudaSetDevice(e->device));
    begin_frame(e, seq);
    SvoSeqCall* hc = next_call_slot(e);
    fill_call(e, hc[0], seq);
    ECUDA(e, upload_left(e, hc[0], img, pitch, false));
    ECUDA(e, push_and_ingest(e, 1));
    ECUDA(e, launch_pyramid(e, dev_call_slot(e), 1, e->st));
    e->launches += e->lay.n_levels - 1;
    // single-sequence view of the [n][max_features] arrays
    int32_t s = seq, c = n;
    return svo_tracker_set(e, 1, &s, feats, ids, &c);
}

int svo_tracker_track(svo_engine* e, int32_t seq, const uint8_t* img, int32_t pitch,
                      int32_t flow_back, float* av_parallax, float* percent_lost) {
    int rc = check_seq(e, seq);
    if (rc) return rc;
    if (!img || pitch < e->p.width || !av_parallax || !percent_lost)
        return fail(e, SVO_ERR_INVALID, "svo_tracker_track: bad arguments");
    SeqState& s = e->seq[seq];
    if (!s.tracker_init) return fail(e, SVO_ERR_STATE, "svo_tracker_track: tracker not initialised");
    ECUDA(e, cudaSetDevice(e->device));
    begin_frame(e, seq);
    SvoSeqCall* hc = next_call_slot(e);
    fill_call(e, hc[0], seq);
    ECUDA(e, cudaEventRecord(e->ev[0], e->st));
    ECUDA(e, upload_left(e, hc[0], img, pitch, false));
    ECUDA(e, cudaEventRecord(e->ev[1], e->st));
    ECUDA(e, push_and_ingest(e, 1));
    ECUDA(e, launch_pyramid(e, dev_call_slot(e), 1, e->st));
    e->launches += e->lay.n_levels - 1;
    ECUDA(e, launch_track(e, 1, s.n_feat, flow_back ? 1 : 0, false));
    ECUDA(e, cudaEventRecord(e->ev[2], e->st));
    SvoTrackStats st;
    ECUDA(e, cudaMemcpyAsync(&st, e->d_stats + seq, sizeof st, cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaEventRecord(e->ev[3], e->st));
    ECUDA(e, wait_stream(e));
    finish_timing(e);
    *av_parallax = st.av_parallax;
    *percent_lost = st.percent_lost;
    s.dbg_attempted = st.n_attempted;
    s.feat_buf = 1 - s.feat_buf;
    s.n_feat = st.n_kept;
    s.last_slot = s.cur_slot;
    return SVO_OK;
}

int svo_tracker_get(svo_engine* e, int32_t seq, svo_float2* feats, uint64_t* ids, int32_t* n) {
    int rc = check_seq(e, seq);
    if (rc) return rc;
    if (!n) return fail(e, SVO_ERR_INVALID, "svo_tracker_get: bad arguments");
    ECUDA(e, cudaSetDevice(e->device));
    const SeqState& s = e->seq[seq];
    const size_t fb = ((size_t)seq * 2 + s.feat_buf) * e->p.max_features;
    *n = s.n_feat;
    if (s.n_feat > 0) {
        if (feats) ECUDA(e, cudaMemcpyAsync(feats, e->d_feat + fb, (size_t)s.n_feat * sizeof(float2), cudaMemcpyDeviceToHost, e->st));
        if (ids) ECUDA(e, cudaMemcpyAsync(ids, e->d_ids + fb, (size_t)s.n_feat * sizeof(uint64_t), cudaMemcpyDeviceToHost, e->st));
        ECUDA(e, cudaStreamSynchronize(e->st));
    }
    return SVO_OK;
}

int svo_tracker_debug(svo_engine* e, int32_t seq, svo_float2* tracked, uint8_t* status1,
                      svo_float2* reverse, uint8_t* status2, int32_t* kept_idx, int32_t* n_attempted) {
    int rc = check_seq(e, seq);
    if (rc) return rc;
    ECUDA(e, cudaSetDevice(e->device));
    const SeqState& s = e->seq[seq];
    const int na = s.dbg_attempted;
    const size_t b = (size_t)seq * e->p.max_features;
    if (n_attempted) *n_attempted = na;
    if (na > 0) {
        if (tracked) ECUDA(e, cudaMemcpyAsync(tracked, e->d_tracked + b, (size_t)na * sizeof(float2), cudaMemcpyDeviceToHost, e->st));
        if (status1) ECUDA(e, cudaMemcpyAsync(status1, e->d_status1 + b, (size_t)na, cudaMemcpyDeviceToHost, e->st));
        if (reverse) ECUDA(e, cudaMemcpyAsync(reverse, e->d_reverse + b, (size_t)na * sizeof(float2), cudaMemcpyDeviceToHost, e->st));
        if (status2) ECUDA(e, cudaMemcpyAsync(status2, e->d_status2 + b, (size_t)na, cudaMemcpyDeviceToHost, e->st));
        ECUDA(e, join_rekey(e));
        if (kept_idx && s.n_feat > 0) ECUDA(e, cudaMemcpyAsync(kept_idx, e->d_kept_idx + b, (size_t)s.n_feat * sizeof(int), cudaMemcpyDeviceToHost, e->st));
        ECUDA(e, cudaStreamSynchronize(e->st));
    }
    return SVO_OK;
}

int svo_dedupe(svo_engine* e, int32_t seq, const svo_float2* det, int32_t n_det,
               const svo_float2* tracked, int32_t n_tracked, float min_distance,
               svo_float2* out_xy, int32_t* out_keep_idx, int32_t* out_n) {
    int rc = check_seq(e, seq);
    if (rc) return rc;
    if (n_det < 0 || n_tracked < 0 || !out_n || (n_det > 0 && !det) || (n_tracked > 0 && !tracked))
        return fail(e, SVO_ERR_INVALID, "svo_dedupe: bad arguments");
    if (n_det > e->pts_cap || n_tracked > e->p.max_features)
        return fail(e, SVO_ERR_CAPACITY, "svo_dedupe: too many points");
    ECUDA(e, cudaSetDevice(e->device));
    const int cap = e->pts_cap, mf = e->p.max_features;
    SvoSeqCall* hc = next_call_slot(e);
    fill_call(e, hc[0], seq);
    hc[0].n_pts = n_det;
    hc[0].n_aux = n_tracked;
    ECUDA(e, push_calls(e, 1));
    // detections go through d_valid2d (scratch), tracked through d_inl
    if (n_det > 0) ECUDA(e, cudaMemcpyAsync(e->d_valid2d + (size_t)seq * cap, det, (size_t)n_det * sizeof(float2), cudaMemcpyHostToDevice, e->st));
    if (n_tracked > 0) ECUDA(e, cudaMemcpyAsync(e->d_inl + (size_t)seq * mf, tracked, (size_t)n_tracked * sizeof(float2), cudaMemcpyHostToDevice, e->st));
    ECUDA(e, svo_launch_dedupe(e->d_valid2d, nullptr, cap, e->d_inl, mf, 0, min_distance, e->d_pts,
                               e->d_keep_idx, e->d_nnew, cap, dev_call_slot(e), e->d_ticket, 1, e->st));
    e->launches += 1;
    int nn = 0;
    ECUDA(e, cudaMemcpyAsync(&nn, e->d_nnew + seq, sizeof(int), cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaStreamSynchronize(e->st));
    *out_n = nn;
    if (nn > 0) {
        if (out_xy) ECUDA(e, cudaMemcpyAsync(out_xy, e->d_pts + (size_t)seq * cap, (size_t)nn * sizeof(float2), cudaMemcpyDeviceToHost, e->st));
        if (out_keep_idx) ECUDA(e, cudaMemcpyAsync(out_keep_idx, e->d_keep_idx + (size_t)seq * cap, (size_t)nn * sizeof(int), cudaMemcpyDeviceToHost, e->st));
        ECUDA(e, cudaStreamSynchronize(e->st));
    }
    return SVO_OK;
}

int svo_debug_min_eigen_map(svo_engine* e, int32_t seq, float* out) {
    int rc = check_seq(e, seq);
    if (rc) return rc;
    if (!out) return fail(e, SVO_ERR_INVALID, "null output");
    if (!e->d_eig) return fail(e, SVO_ERR_STATE, "eig map not kept: create the engine with keep_eig_map = 1");
    ECUDA(e, cudaSetDevice(e->device));
    ECUDA(e, cudaMemcpy2DAsync(out, (size_t)e->p.width * sizeof(float), e->d_eig + (size_t)seq * e->eig_stride,
                               (size_t)e->eig_pitch * sizeof(float), (size_t)e->p.width * sizeof(float),
                               e->p.height, cudaMemcpyDeviceToHost, e->st));
    ECUDA(e, cudaStreamSynchronize(e->st));
    return SVO_OK;
}

int svo_debug_pyramid_level(svo_engine* e, int32_t seq, int32_t level, uint8_t* out, int32_t* w, int32_t* h) {
    int rc = check_seq(e, seq);
    if (rc) return rc;
    if (level < 0 || level >= e->lay.n_levels) return fail(e, SVO_ERR_INVALID, "level out of range");
    const SvoLevel& lv = e->lay.lvl[level];
    if (w) *w = lv.w;
    if (h) *h = lv.h;
    if (out) {
        ECUDA(e, cudaSetDevice(e->device));
        ECUDA(e, cudaMemcpy2DAsync(out, lv.w, slot_ptr(e, seq, e->seq[seq].cur_slot) + lv.off, lv.pitch,
                                   lv.w, lv.h, cudaMemcpyDeviceToHost, e->st));
        ECUDA(e, cudaStreamSynchronize(e->st));
    }
    return SVO_OK;
}

int svo_debug_candidates(svo_engine* e, int32_t seq, int32_t* xy, int32_t cap, int32_t* n) {
    int rc = check_seq(e, seq);
    if (rc) return rc;
    if (!n) return fail(e, SVO_ERR_INVALID, "null output");
    ECUDA(e, cudaSetDevice(e->device));
    int cnt = 0;
    ECUDA(e, cudaMemcpy(&cnt, e->d_cand_count + seq, sizeof(int), cudaMemcpyDeviceToHost));
    const unsigned long long* src = e->d_cand + (size_t)seq * e->max_cand;
    if (cnt > e->max_cand) {
        // the sequence's last detection went through the spill lists (one sequence at a time)
        if (e->spill_seq != seq || cnt > e->max_cand_big)
            return fail(e, SVO_ERR_STATE, "svo_debug_candidates: the spilled candidate list of this sequence was reused by another sequence");
        src = e->d_cand_big;
    }
    *n = cnt;
    int m = cnt < cap ? cnt : cap;
    if (xy && m > 0) {
        std::vector<unsigned long long> keys(m);
        ECUDA(e, cudaMemcpy(keys.data(), src, (size_t)m * 8, cudaMemcpyDeviceToHost));
        for (int i = 0; i < m; ++i) {
            unsigned addr = (unsigned)(keys[i] & 0xffffffffull);
            xy[2 * i] = (int)(addr % (unsigned)e->p.width);
            xy[2 * i + 1] = (int)(addr / (unsigned)e->p.width);
        }
    }
    return SVO_OK;
}

int svo_device_alloc(svo_engine* e, size_t bytes, void** out) {
    if (!e || !out) return SVO_ERR_INVALID;
    ECUDA(e, cudaSetDevice(e->device));
    ECUDA(e, cudaMalloc(out, bytes));
    return SVO_OK;
}
int svo_device_free(svo_engine* e, void* p) {
    if (!e) return SVO_ERR_INVALID;
    ECUDA(e, cudaSetDevice(e->device));
    ECUDA(e, cudaFree(p));
    return SVO_OK;
}
int svo_device_upload(svo_engine* e, void* dst, const void* src, size_t bytes) {
    if (!e) return SVO_ERR_INVALID;
    ECUDA(e, cudaSetDevice(e->device));
    ECUDA(e, cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
    return SVO_OK;
}
int svo_host_alloc(size_t bytes, void** out) {
    if (!out) return SVO_ERR_INVALID;
    // mapped + portable: kernels of any device may read the buffer in place (svo_device_can_read)
    return cudaHostAlloc(out, bytes, cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess
               ? SVO_OK : SVO_ERR_NOMEM;
}
int svo_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? SVO_OK : SVO_ERR_CUDA; }

}  // extern "C"
