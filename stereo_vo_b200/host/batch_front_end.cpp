// BatchFrontEnd: S independent sequences on one GPU, G engine groups, one host worker per group
// (see include/stereo_vo/batch_front_end.hpp). All pixel work is inside libsvo.so.
#include "stereo_vo/batch_front_end.hpp"

#include <algorithm>
#include <cstring>
#include <stdexcept>

namespace svo {

namespace {
std::string engine_error(svo_engine* e, const char* what, int rc) {
    const char* msg = svo_last_error(e);
    return std::string(what) + " failed (" + std::to_string(rc) + "): " + (msg ? msg : "");
}
}  // namespace

BatchFrontEnd::BatchFrontEnd(const BatchOptions& opt) : opt_(opt) {
    if (opt.n_seq <= 0) throw std::invalid_argument("BatchFrontEnd: n_seq must be positive");
    const int G = std::max(1, std::min(opt.n_groups, opt.n_seq));
    const int mc = opt.params.max_corners, mf = opt.params.max_features;
    seq_.resize((size_t)opt.n_seq);
    corners_.resize((size_t)opt.n_seq * mc);
    tracker_pts_.resize((size_t)opt.n_seq * mf);
    tracker_ids_.resize((size_t)opt.n_seq * mf);
    new_2d_.resize((size_t)opt.n_seq * mc);
    new_3d_.resize((size_t)opt.n_seq * mc);
    // contiguous blocks of sequences per group (sizes differ by at most one)
    int s0 = 0;
    for (int g = 0; g < G; ++g) {
        const int cnt = opt.n_seq / G + (g < opt.n_seq % G ? 1 : 0);
        auto grp = std::make_unique<Group>();
        int rc = svo_create(&opt.params, opt.device, cnt, &grp->e);
        if (rc != SVO_OK) {
            std::string msg = engine_error(nullptr, "svo_create", rc);
            for (auto& done : groups_) svo_destroy(done->e);
            throw std::runtime_error(msg);
        }
        for (int i = 0; i < cnt; ++i) {
            grp->seqs.push_back(s0 + i);
            seq_[(size_t)(s0 + i)].group = g;
            seq_[(size_t)(s0 + i)].local = i;
        }
        s0 += cnt;
        const size_t n = (size_t)cnt;
        grp->local_ids.resize(n); grp->kf_local.resize(n); grp->n_inl.resize(n);
        grp->n_newv.resize(n); grp->n_valid.resize(n); grp->counts.resize(n); grp->n_corn.resize(n);
        grp->lp.resize(n); grp->rp.resize(n); grp->nlp.resize(n); grp->nrp.resize(n);
        grp->stats.resize(n);
        grp->corners.resize(n * mc); grp->new_feats.resize(n * mc); grp->valid_2d.resize(n * mc);
        grp->pts_3d.resize(n * mc); grp->valid_idx.resize(n * mc); grp->disp16.resize(n * mc);
        grp->tracked.resize(n * mf); grp->tracked_ids.resize(n * mf); grp->inl.resize(n * mf);
        grp->feats.resize(n * mf); grp->ids.resize(n * mf);
        grp->poses.resize(n * 16);
        for (int i = 0; i < cnt; ++i) grp->local_ids[(size_t)i] = i;
        groups_.push_back(std::move(grp));
    }
    if (G > 1)
        for (auto& g : groups_) g->worker = std::thread(&BatchFrontEnd::worker_loop, this, g.get());
}

BatchFrontEnd::~BatchFrontEnd() {
    {
        std::lock_guard<std::mutex> lk(mu_);
        stop_ = true;
    }
    cv_start_.notify_all();
    for (auto& g : groups_) {
        if (g->worker.joinable()) g->worker.join();
        if (g->e) svo_destroy(g->e);
    }
}

void BatchFrontEnd::worker_loop(Group* g) {
    uint64_t seen = 0;
    for (;;) {
        {
            std::unique_lock<std::mutex> lk(mu_);
            cv_start_.wait(lk, [&] { return stop_ || generation_ != seen; });
            if (stop_) return;
            seen = generation_;
        }
        try {
            run_group(*g);
        } catch (const std::exception& ex) {
            g->error = ex.what();
        }
        {
            std::lock_guard<std::mutex> lk(mu_);
            --pending_;
        }
        cv_done_.notify_one();
    }
}

// One frame for the sequences of one group: the body of ImageProcessor::process, batched.
void BatchFrontEnd::run_group(Group& g) {
    const Job& job = job_;
    svo_engine* e = g.e;
    const int n = (int)g.seqs.size();
    const int mc = opt_.params.max_corners, mf = opt_.params.max_features;
    g.sum = BatchSummary();
    g.error.clear();
    for (int i = 0; i < n; ++i) {
        const int s = g.seqs[(size_t)i];
        g.lp[(size_t)i] = job.left[s];
        g.rp[(size_t)i] = job.right ? job.right[s] : nullptr;
    }
    int rc;
    if (!job.on_device && job.next_left) {
        // announce the next step's frames first: their copies run on the engine's copy stream
        // underneath the kernels launched below
        for (int i = 0; i < n; ++i) {
            const int s = g.seqs[(size_t)i];
            g.nlp[(size_t)i] = job.next_left[s];
            g.nrp[(size_t)i] = job.next_right ? job.next_right[s] : nullptr;
        }
        rc = svo_prefetch(e, n, g.local_ids.data(), g.nlp.data(), g.nrp.data(), job.pitch);
        if (rc != SVO_OK) throw std::runtime_error(engine_error(e, "svo_prefetch", rc));
    }
    rc = job.on_device
        ? svo_step_device(e, n, g.local_ids.data(), g.lp.data(), g.rp.data(), job.pitch,
                          g.corners.data(), g.n_corn.data(), g.stats.data(), g.tracked.data(),
                          g.tracked_ids.data())
        : svo_step(e, n, g.local_ids.data(), g.lp.data(), g.rp.data(), job.pitch, g.corners.data(),
                   g.n_corn.data(), g.stats.data(), g.tracked.data(), g.tracked_ids.data());
    if (rc != SVO_OK) throw std::runtime_error(engine_error(e, "svo_step", rc));

    // ---- keyframe decision per sequence (src/image_processor.cpp:23-30,63-65)
    int nk = 0;
    for (int i = 0; i < n; ++i) {
        const int s = g.seqs[(size_t)i];
        SeqState& st = seq_[(size_t)s];
        st.n_corners = g.n_corn[(size_t)i];
        st.stats = g.stats[(size_t)i];
        st.n_new = 0;
        std::memcpy(&corners_[(size_t)s * mc], &g.corners[(size_t)i * mc],
                    (size_t)st.n_corners * sizeof(svo_float2));
        g.sum.frames += 1;
        g.sum.corners += st.n_corners;
        if (st.n_corners < 4) {
            st.decision = BATCH_TOO_FEW;
            continue;
        }
        int n_inl = 0;
        if (st.have_kf) {
            const svo_track_stats& ts = st.stats;
            st.n_tracker = ts.n_kept;
            std::memcpy(&tracker_pts_[(size_t)s * mf], &g.tracked[(size_t)i * mf],
                        (size_t)ts.n_kept * sizeof(svo_float2));
            std::memcpy(&tracker_ids_[(size_t)s * mf], &g.tracked_ids[(size_t)i * mf],
                        (size_t)ts.n_kept * sizeof(uint64_t));
            g.sum.tracked += ts.n_kept;
            if (ts.av_parallax <= opt_.parallax_thresh && ts.percent_lost < 0.4) {
                st.decision = BATCH_SKIP;
                continue;
            }
            st.decision = BATCH_KEYFRAME;
            n_inl = ts.n_kept;             // every kept track is an inlier (no host PnP here)
            std::memcpy(&g.inl[(size_t)nk * mf], &g.tracked[(size_t)i * mf],
                        (size_t)n_inl * sizeof(svo_float2));
        } else {
            st.decision = BATCH_FIRST;
        }
        g.kf_local[(size_t)nk] = i;
        g.n_inl[(size_t)nk] = n_inl;
        float* pose = &g.poses[(size_t)nk * 16];
        if (job.poses) std::memcpy(pose, job.poses + (size_t)s * 16, 16 * sizeof(float));
        else for (int k = 0; k < 16; ++k) pose[k] = (k % 5 == 0) ? 1.f : 0.f;
        ++nk;
    }
    if (nk == 0) return;

    // ---- dedupe + StereoBM + reprojection for the new keyframes (:113-142)
    rc = svo_keyframe(e, nk, g.kf_local.data(), g.inl.data(), g.n_inl.data(), g.poses.data(),
                      opt_.focal, opt_.cx, opt_.cy, opt_.baseline, g.new_feats.data(),
                      g.n_newv.data(), g.valid_2d.data(), g.pts_3d.data(), g.valid_idx.data(),
                      g.disp16.data(), g.n_valid.data());
    if (rc != SVO_OK) throw std::runtime_error(engine_error(e, "svo_keyframe", rc));

    // ---- BundleAdjuster::add_keyframe bookkeeping (ids are a running counter, SURVEY.md B7;
    // new features are clipped to the 400-feature limit) + FeatureTracker::init (:148-162)
    for (int k = 0; k < nk; ++k) {
        const int i = g.kf_local[(size_t)k];
        const int s = g.seqs[(size_t)i];
        SeqState& st = seq_[(size_t)s];
        const int n_inl = g.n_inl[(size_t)k];
        const int nv = std::min(g.n_valid[(size_t)k], mf - n_inl);
        svo_float2* f = &g.feats[(size_t)k * mf];
        uint64_t* id = &g.ids[(size_t)k * mf];
        std::memcpy(f, &g.inl[(size_t)k * mf], (size_t)n_inl * sizeof(svo_float2));
        std::memcpy(id, &g.tracked_ids[(size_t)i * mf], (size_t)n_inl * sizeof(uint64_t));
        std::memcpy(f + n_inl, &g.valid_2d[(size_t)k * mc], (size_t)nv * sizeof(svo_float2));
        for (int j = 0; j < nv; ++j) id[n_inl + j] = st.next_id++;
        g.counts[(size_t)k] = n_inl + nv;
        st.have_kf = true;
        st.n_new = nv;
        st.n_tracker = n_inl + nv;
        std::memcpy(&new_2d_[(size_t)s * mc], &g.valid_2d[(size_t)k * mc], (size_t)nv * sizeof(svo_float2));
        std::memcpy(&new_3d_[(size_t)s * mc], &g.pts_3d[(size_t)k * mc], (size_t)nv * sizeof(svo_float3));
        std::memcpy(&tracker_pts_[(size_t)s * mf], f, (size_t)(n_inl + nv) * sizeof(svo_float2));
        std::memcpy(&tracker_ids_[(size_t)s * mf], id, (size_t)(n_inl + nv) * sizeof(uint64_t));
        g.sum.keyframes += 1;
        g.sum.new_points += nv;
    }
    rc = svo_tracker_set(e, nk, g.kf_local.data(), g.feats.data(), g.ids.data(), g.counts.data());
    if (rc != SVO_OK) throw std::runtime_error(engine_error(e, "svo_tracker_set", rc));
}

BatchSummary BatchFrontEnd::step(const uint8_t* const* left, const uint8_t* const* right, int pitch,
                                 bool on_device, const float* poses,
                                 const uint8_t* const* next_left,
                                 const uint8_t* const* next_right) {
    if (!left) throw std::invalid_argument("BatchFrontEnd::step: left is null");
    job_.left = left; job_.right = right; job_.pitch = pitch; job_.on_device = on_device;
    job_.poses = poses; job_.next_left = next_left; job_.next_right = next_right;
    if (groups_.size() == 1) {
        run_group(*groups_[0]);
    } else {
        {
            std::lock_guard<std::mutex> lk(mu_);
            pending_ = (int)groups_.size();
            ++generation_;
        }
        cv_start_.notify_all();
        std::unique_lock<std::mutex> lk(mu_);
        cv_done_.wait(lk, [&] { return pending_ == 0; });
    }
    BatchSummary total;
    for (auto& g : groups_) {
        if (!g->error.empty()) throw std::runtime_error(g->error);
        total.frames += g->sum.frames; total.keyframes += g->sum.keyframes;
        total.corners += g->sum.corners; total.tracked += g->sum.tracked;
        total.new_points += g->sum.new_points;
    }
    return total;
}

const svo_float2* BatchFrontEnd::corners(int s) const { return &corners_[(size_t)s * opt_.params.max_corners]; }
const svo_float2* BatchFrontEnd::tracker_points(int s) const { return &tracker_pts_[(size_t)s * opt_.params.max_features]; }
const uint64_t* BatchFrontEnd::tracker_ids(int s) const { return &tracker_ids_[(size_t)s * opt_.params.max_features]; }
const svo_float2* BatchFrontEnd::new_2d(int s) const { return &new_2d_[(size_t)s * opt_.params.max_corners]; }
const svo_float3* BatchFrontEnd::new_3d(int s) const { return &new_3d_[(size_t)s * opt_.params.max_corners]; }

int64_t BatchFrontEnd::kernel_launches() const {
    int64_t n = 0;
    for (auto& g : groups_) n += svo_kernel_launches(g->e);
    return n;
}

void BatchFrontEnd::timer_begin() {
    for (auto& g : groups_) svo_timer_begin(g->e);
}

float BatchFrontEnd::timer_end() {
    float worst = 0.f;
    for (auto& g : groups_) {
        float ms = 0.f;
        svo_timer_end(g->e, &ms);
        worst = std::max(worst, ms);
    }
    return worst;
}

void BatchFrontEnd::set_profiling(bool on) {
    for (auto& g : groups_) svo_set_profiling(g->e, on ? 1 : 0);
}

void BatchFrontEnd::get_profile(std::vector<double>& ms, std::vector<int64_t>& counts) const {
    const int slots = svo_profile_slots();
    ms.assign((size_t)slots, 0.0);
    counts.assign((size_t)slots, 0);
    std::vector<double> m((size_t)slots);
    std::vector<int64_t> c((size_t)slots);
    for (auto& g : groups_) {
        svo_get_profile(g->e, m.data(), c.data());
        for (int i = 0; i < slots; ++i) { ms[(size_t)i] += m[(size_t)i]; counts[(size_t)i] += c[(size_t)i]; }
    }
}

}  // namespace svo
