// BatchFrontEnd: S independent sequences on one GPU, G engine groups, one host worker per group
// (see include/stereo_vo/batch_front_end.hpp). All pixel work is inside libsvo.so.
#include "stereo_vo/batch_front_end.hpp"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
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
    if (const char* v = std::getenv("SVO_BATCH_LAZY_RIGHT")) lazy_right_ = std::atoi(v) != 0;
    if (const char* v = std::getenv("SVO_BATCH_TRACE")) trace_ = std::atoi(v) != 0;
    if (const char* v = std::getenv("SVO_BATCH_ZERO_COPY_RIGHT")) zero_copy_right_ = std::atoi(v) != 0;
    if (opt.n_seq <= 0) throw std::invalid_argument("BatchFrontEnd: n_seq must be positive");
    const int G = std::max(1, std::min(opt.n_groups, opt.n_seq));
    const int mc = opt.params.max_corners, mf = opt.params.max_features;
    seq_.resize((size_t)opt.n_seq);
    for (int p = 0; p < kPipeDepth; ++p) {
        corners_[p].resize((size_t)opt.n_seq * mc);
        tracker_pts_[p].resize((size_t)opt.n_seq * mf);
        tracker_ids_[p].resize((size_t)opt.n_seq * mf);
        new_2d_[p].resize((size_t)opt.n_seq * mc);
        new_3d_[p].resize((size_t)opt.n_seq * mc);
        new_feats_[p].resize((size_t)opt.n_seq * mc);
        disp16_[p].resize((size_t)opt.n_seq * mc);
        valid_idx_[p].resize((size_t)opt.n_seq * mc);
        jobs_[p].sum.resize((size_t)G);
        jobs_[p].error.resize((size_t)G);
    }
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
        grp->srp.resize(n); grp->miss_local.resize(n);
        grp->pred_cur.assign(n, 0); grp->pred_next.assign(n, 0);
        grp->stats.resize(n);
        grp->corners.resize(n * mc); grp->new_feats.resize(n * mc); grp->valid_2d.resize(n * mc);
        grp->pts_3d.resize(n * mc); grp->valid_idx.resize(n * mc); grp->disp16.resize(n * mc);
        grp->tracked.resize(n * mf); grp->tracked_ids.resize(n * mf); grp->inl.resize(n * mf);
        grp->feats.resize(n * mf); grp->ids.resize(n * mf); grp->first_id.resize(n);
        grp->poses.resize(n * 16);
        grp->pnp_world.resize(n * mf); grp->pnp_image.resize(n * mf); grp->pnp_inl.resize(n * mf);
        grp->pnp_counts.resize(n); grp->pnp_kf.resize(n); grp->pnp_guess.resize(n * 6);
        grp->pnp_res.resize(n);
        for (int i = 0; i < cnt; ++i) grp->local_ids[(size_t)i] = i;
        groups_.push_back(std::move(grp));
    }
    int gi = 0;
    for (auto& g : groups_) g->worker = std::thread(&BatchFrontEnd::worker_loop, this, g.get(), gi++);
}

BatchFrontEnd::~BatchFrontEnd() {
    {
        std::lock_guard<std::mutex> lk(mu_);
        stop_ = true;
    }
    cv_start_.notify_all();
    int gi = 0;
    for (auto& g : groups_) {
        if (g->worker.joinable()) g->worker.join();
        if (trace_ && g->phase_calls) {
            const double c = (double)g->phase_calls;
            std::fprintf(stderr,
                         "[svo batch] group %d (%zu seqs, %ld steps) host ms/step: prefetch %.3f "
                         "step %.3f set_right %.3f keyframe %.3f bookkeeping %.3f tracker_set %.3f | "
                         "device ms/step: step copy-in %.3f kernels %.3f copy-out %.3f ; keyframe "
                         "copy-in %.3f kernels %.3f copy-out %.3f | right images: %ld sent ahead, %ld "
                         "keyframes predicted, %ld fetched on demand\n",
                         gi, g->seqs.size(), (long)g->phase_calls, g->phase_ms[0] / c,
                         g->phase_ms[1] / c, g->phase_ms[2] / c, g->phase_ms[3] / c,
                         g->phase_ms[4] / c, g->phase_ms[5] / c, g->phase_ms[6] / c,
                         g->phase_ms[7] / c, g->phase_ms[8] / c, g->phase_ms[9] / c,
                         g->phase_ms[10] / c, g->phase_ms[11] / c, (long)g->pred_sent,
                         (long)g->pred_hits, (long)g->pred_miss);
        }
        ++gi;
        if (g->e) svo_destroy(g->e);
    }
}

void BatchFrontEnd::worker_loop(Group* g, int gi) {
    for (;;) {
        uint64_t job_no;
        {
            std::unique_lock<std::mutex> lk(mu_);
            cv_start_.wait(lk, [&] { return stop_ || n_submitted_ > g->done; });
            if (stop_) return;
            job_no = g->done;
        }
        const int slot = (int)(job_no % kPipeDepth);
        Job& job = jobs_[slot];
        try {
            run_group(*g, job, slot, gi);
        } catch (const std::exception& ex) {
            job.error[(size_t)gi] = ex.what();
        }
        {
            std::lock_guard<std::mutex> lk(mu_);
            g->done = job_no + 1;
            --job.pending;
        }
        cv_done_.notify_all();
    }
}

namespace {
inline double now_ms() {
    return std::chrono::duration<double, std::milli>(
               std::chrono::steady_clock::now().time_since_epoch()).count();
}
}  // namespace

// One frame for the sequences of one group: the body of ImageProcessor::process, batched.
void BatchFrontEnd::run_group(Group& g, Job& job, int slot, int gi) {
    double t_mark = trace_ ? now_ms() : 0.0;
    auto lap = [&](int slot) {          // host wall time per phase (SVO_BATCH_TRACE=1)
        if (!trace_) return;
        const double t = now_ms();
        g.phase_ms[slot] += t - t_mark;
        t_mark = t;
    };
    g.phase_calls += 1;
    svo_engine* e = g.e;
    BatchSummary& sum = job.sum[(size_t)gi];
    sum = BatchSummary();
    job.error[(size_t)gi].clear();
    const int n = (int)g.seqs.size();
    const int mc = opt_.params.max_corners, mf = opt_.params.max_features;
    for (int i = 0; i < n; ++i) {
        const int s = g.seqs[(size_t)i];
        g.lp[(size_t)i] = job.left[s];
        g.rp[(size_t)i] = job.has_right ? job.right[s] : nullptr;
    }
    // The right image is only read by the stereo stage of a keyframe
    // (src/image_processor.cpp:34-39,137-142). With lazy_right_ it travels with the step only for
    // the sequences PREDICTED to become keyframes (parallax extrapolated from the last steps:
    // a correct guess keeps the transfer off this group's critical chain); a keyframe that was
    // not predicted fetches its right image on demand (svo_set_right), a wrong guess costs one
    // idle image copy. Results never depend on the prediction.
    const bool lazy = lazy_right_ && job.has_right;
    auto predict = [&](const SeqState& st, int steps_ahead) {
        if (!st.have_kf) return steps_ahead == 1;                     // the bootstrap keyframe
        const float margin = 0.85f * opt_.parallax_thresh;
        const float p1 = st.par_last + st.par_rate;                   // next step
        if (steps_ahead == 1) return p1 > margin;
        if (p1 > opt_.parallax_thresh) return st.par_rate > margin;   // keyframe resets parallax
        return st.par_last + 2.f * st.par_rate > margin;
    };
    if (lazy) {
        // right images that travel with THIS step: what the previous job announced, else a
        // fresh one-step prediction
        if (!g.have_pred) {
            for (int i = 0; i < n; ++i) g.pred_cur[(size_t)i] = predict(seq_[(size_t)g.seqs[(size_t)i]], 1);
        } else {
            g.pred_cur = g.pred_next;
        }
        g.have_pred = false;
        for (int i = 0; i < n; ++i)
            g.srp[(size_t)i] = g.pred_cur[(size_t)i] ? g.rp[(size_t)i] : nullptr;
    }
    const uint8_t* const* step_right = lazy ? g.srp.data() : g.rp.data();
    int rc;
    if (!job.on_device && job.has_next) {
        // announce the next step's frames first: their copies run on the engine's copy stream
        // underneath the kernels launched below
        for (int i = 0; i < n; ++i) {
            const int s = g.seqs[(size_t)i];
            g.nlp[(size_t)i] = job.next_left[s];
            bool want_right = job.has_next_right;
            if (lazy && want_right) {
                want_right = predict(seq_[(size_t)s], 2);
                g.pred_next[(size_t)i] = want_right;
            }
            g.nrp[(size_t)i] = want_right ? job.next_right[s] : nullptr;
        }
        if (lazy && job.has_next_right) g.have_pred = true;
        rc = svo_prefetch(e, n, g.local_ids.data(), g.nlp.data(), g.nrp.data(), job.pitch);
        if (rc != SVO_OK) throw std::runtime_error(engine_error(e, "svo_prefetch", rc));
    }
    lap(0);
    rc = job.on_device
        ? svo_step_device(e, n, g.local_ids.data(), g.lp.data(), step_right, job.pitch,
                          g.corners.data(), g.n_corn.data(), g.stats.data(), g.tracked.data(),
                          g.tracked_ids.data())
        : svo_step(e, n, g.local_ids.data(), g.lp.data(), step_right, job.pitch, g.corners.data(),
                   g.n_corn.data(), g.stats.data(), g.tracked.data(), g.tracked_ids.data());
    if (rc != SVO_OK) throw std::runtime_error(engine_error(e, "svo_step", rc));
    lap(1);
    if (trace_) {
        float ms[3];
        svo_last_timing(e, ms);
        g.phase_ms[6] += ms[0]; g.phase_ms[7] += ms[1]; g.phase_ms[8] += ms[2];
    }

    // ---- keyframe decision per sequence (src/image_processor.cpp:23-30,63-65)
    int nk = 0;
    for (int i = 0; i < n; ++i) {
        const int s = g.seqs[(size_t)i];
        SeqState& st = seq_[(size_t)s];
        SeqResult& rs = st.res[slot];
        rs.n_corners = g.n_corn[(size_t)i];
        rs.stats = g.stats[(size_t)i];
        rs.n_new = 0;
        rs.n_new_features = rs.n_valid = rs.n_inliers = 0;
        rs.n_tracker = st.n_tracker;
        std::memcpy(&corners_[slot][(size_t)s * mc], &g.corners[(size_t)i * mc],
                    (size_t)rs.n_corners * sizeof(svo_float2));
        sum.frames += 1;
        sum.corners += rs.n_corners;
        if (rs.n_corners < 4) {
            rs.decision = BATCH_TOO_FEW;
            continue;
        }
        int n_inl = 0;
        if (st.have_kf) {
            const svo_track_stats& ts = rs.stats;
            st.n_tracker = rs.n_tracker = ts.n_kept;
            std::memcpy(&tracker_pts_[slot][(size_t)s * mf], &g.tracked[(size_t)i * mf],
                        (size_t)ts.n_kept * sizeof(svo_float2));
            std::memcpy(&tracker_ids_[slot][(size_t)s * mf], &g.tracked_ids[(size_t)i * mf],
                        (size_t)ts.n_kept * sizeof(uint64_t));
            sum.tracked += ts.n_kept;
            // parallax model of the keyframe predictor (finite values only)
            if (ts.av_parallax == ts.av_parallax) {
                st.par_rate = 0.5f * st.par_rate + 0.5f * std::max(0.f, ts.av_parallax - st.par_last);
                st.par_last = ts.av_parallax;
            }
            if (ts.av_parallax <= opt_.parallax_thresh && ts.percent_lost < 0.4) {
                rs.decision = BATCH_SKIP;
                continue;
            }
            rs.decision = BATCH_KEYFRAME;
            st.par_last = 0.f;             // the tracker restarts from this frame
            n_inl = ts.n_kept;             // every kept track is an inlier (no host PnP here)
            std::memcpy(&g.inl[(size_t)nk * mf], &g.tracked[(size_t)i * mf],
                        (size_t)n_inl * sizeof(svo_float2));
        } else {
            rs.decision = BATCH_FIRST;
        }
        g.kf_local[(size_t)nk] = i;
        g.n_inl[(size_t)nk] = n_inl;
        float* pose = &g.poses[(size_t)nk * 16];
        // the first keyframe is triangulated with the identity pose, whatever the caller
        // supplies (src/image_processor.cpp:34-39: cv::Mat::eye(4, 4, CV_32F))
        if (job.has_poses && rs.decision != BATCH_FIRST)
            std::memcpy(pose, job.poses.data() + (size_t)s * 16, 16 * sizeof(float));
        else for (int k = 0; k < 16; ++k) pose[k] = (k % 5 == 0) ? 1.f : 0.f;
        ++nk;
    }
    if (lazy && nk > 0) {
        // keyframes the predictor missed: fetch their right image now. Pinned, device-mapped
        // host frames are read by the ingest kernel itself (zero-copy over PCIe): a DMA copy
        // would queue behind the bulk prefetches of every group on the one H2D copy engine
        int nm = 0;
        for (int k = 0; k < nk; ++k) {
            const int i = g.kf_local[(size_t)k];
            if (g.pred_cur[(size_t)i]) { g.pred_hits += 1; continue; }
            g.miss_local[(size_t)nm] = i;
            g.nrp[(size_t)nm] = g.rp[(size_t)i];
            ++nm;
        }
        g.pred_miss += nm;
        sum.right_images += nm;
        if (nm > 0) {
            bool direct = job.on_device;
            if (!direct && zero_copy_right_) {
                direct = true;
                for (int k = 0; k < nm && direct; ++k) direct = svo_device_can_read(g.nrp[(size_t)k]) != 0;
            }
            rc = svo_set_right(e, nm, g.miss_local.data(), g.nrp.data(), job.pitch, direct ? 1 : 0);
            if (rc != SVO_OK) throw std::runtime_error(engine_error(e, "svo_set_right", rc));
        }
    }
    if (lazy) {
        for (int i = 0; i < n; ++i) {
            g.pred_sent += g.pred_cur[(size_t)i] ? 1 : 0;
            sum.right_images += g.pred_cur[(size_t)i] ? 1 : 0;
        }
    } else if (job.has_right) {
        sum.right_images += n;
    }
    lap(2);
    if (nk == 0) return;
    if (closed_loop_) {
        run_keyframes_closed_loop(g, slot, nk, sum);
        lap(3);
        return;
    }

    // ---- dedupe + StereoBM + reprojection for the new keyframes (:113-142), then
    // BundleAdjuster::add_keyframe's id assignment (a running counter; new features
    // clipped to the 400-feature limit) and FeatureTracker::init (:148-162). Every tracked
    // feature is an inlier here, so the engine takes the inlier list from its own tracker state
    // and re-initialises the tracker on the device (svo_keyframe_tracked): no list goes up.
    for (int k = 0; k < nk; ++k) {
        const int s = g.seqs[(size_t)g.kf_local[(size_t)k]];
        g.first_id[(size_t)k] = seq_[(size_t)s].next_id;
    }
    rc = svo_keyframe_tracked(e, nk, g.kf_local.data(), g.poses.data(), g.first_id.data(),
                              opt_.focal, opt_.cx, opt_.cy, opt_.baseline, g.new_feats.data(),
                              g.n_newv.data(), g.valid_2d.data(), g.pts_3d.data(),
                              g.valid_idx.data(), g.disp16.data(), g.n_valid.data());
    if (rc != SVO_OK) throw std::runtime_error(engine_error(e, "svo_keyframe_tracked", rc));
    lap(3);
    if (trace_) {
        float ms[3];
        svo_last_timing(e, ms);
        g.phase_ms[9] += ms[0]; g.phase_ms[10] += ms[1]; g.phase_ms[11] += ms[2];
    }

    // the host's copy of the new tracker set (results of this step)
    for (int k = 0; k < nk; ++k) {
        const int i = g.kf_local[(size_t)k];
        const int s = g.seqs[(size_t)i];
        SeqState& st = seq_[(size_t)s];
        const int n_inl = g.n_inl[(size_t)k];
        const int nv = std::min(g.n_valid[(size_t)k], mf - n_inl);
        svo_float2* f = &tracker_pts_[slot][(size_t)s * mf];
        uint64_t* id = &tracker_ids_[slot][(size_t)s * mf];
        std::memcpy(f, &g.inl[(size_t)k * mf], (size_t)n_inl * sizeof(svo_float2));
        std::memcpy(id, &g.tracked_ids[(size_t)i * mf], (size_t)n_inl * sizeof(uint64_t));
        std::memcpy(f + n_inl, &g.valid_2d[(size_t)k * mc], (size_t)nv * sizeof(svo_float2));
        for (int j = 0; j < nv; ++j) id[n_inl + j] = st.next_id++;
        st.have_kf = true;
        SeqResult& rs = st.res[slot];
        rs.n_new = nv;
        rs.n_new_features = g.n_newv[(size_t)k];
        rs.n_valid = g.n_valid[(size_t)k];
        std::memcpy(&new_feats_[slot][(size_t)s * mc], &g.new_feats[(size_t)k * mc], (size_t)rs.n_new_features * sizeof(svo_float2));
        std::memcpy(&disp16_[slot][(size_t)s * mc], &g.disp16[(size_t)k * mc], (size_t)rs.n_new_features * sizeof(int16_t));
        std::memcpy(&valid_idx_[slot][(size_t)s * mc], &g.valid_idx[(size_t)k * mc], (size_t)rs.n_valid * sizeof(int32_t));
        st.n_tracker = rs.n_tracker = n_inl + nv;
        std::memcpy(&new_2d_[slot][(size_t)s * mc], &g.valid_2d[(size_t)k * mc], (size_t)nv * sizeof(svo_float2));
        std::memcpy(&new_3d_[slot][(size_t)s * mc], &g.pts_3d[(size_t)k * mc], (size_t)nv * sizeof(svo_float3));
        sum.keyframes += 1;
        sum.new_points += nv;
    }
    lap(4);
}

// Keyframe stage with the pose step in the loop (src/image_processor.cpp:71-162), for the nk
// keyframes of one group: PnP-RANSAC on the GPU, inlier keyframe, dedupe + stereo + reprojection
// with hmat, id assignment, tracker re-init.
void BatchFrontEnd::run_keyframes_closed_loop(Group& g, int slot, int nk, BatchSummary& sum) {
    svo_engine* e = g.e;
    const int mc = opt_.params.max_corners, mf = opt_.params.max_features;
    // ---- pose step for the keyframes that have tracked features (not the first keyframe)
    int np = 0;
    for (int k = 0; k < nk; ++k) {
        const int i = g.kf_local[(size_t)k];
        SeqState& st = seq_[(size_t)g.seqs[(size_t)i]];
        if (st.res[slot].decision != BATCH_KEYFRAME) continue;
        const int n = st.res[slot].stats.n_kept;
        for (int j = 0; j < n; ++j) {
            const uint64_t id = g.tracked_ids[(size_t)i * mf + j];
            if (id >= st.world.size()) throw std::out_of_range("BatchFrontEnd: tracked id without a world point");
            g.pnp_world[(size_t)np * mf + j] = st.world[(size_t)id];    // get_world_points(ids), :72
            g.pnp_image[(size_t)np * mf + j] = g.tracked[(size_t)i * mf + j];
        }
        g.pnp_counts[(size_t)np] = n;
        std::memcpy(&g.pnp_guess[(size_t)np * 6], st.rt, sizeof st.rt);
        g.pnp_kf[(size_t)np] = k;
        ++np;
    }
    if (np > 0) {
        int rc = svo_pnp_ransac(e, np, g.pnp_world.data(), g.pnp_image.data(), g.pnp_counts.data(),
                                opt_.focal, opt_.cx, opt_.cy, pnp_iterations_, pnp_reproj_,
                                pnp_confidence_, g.pnp_guess.data(), g.pnp_res.data(), g.pnp_inl.data());
        if (rc != SVO_OK) throw std::runtime_error(engine_error(e, "svo_pnp_ransac", rc));
    }
    // ---- inlier lists and poses (:95-108, :130-134); a first keyframe has none and the identity
    for (int k = 0; k < nk; ++k) {
        g.n_inl[(size_t)k] = 0;
        float* pose = &g.poses[(size_t)k * 16];
        for (int q = 0; q < 16; ++q) pose[q] = (q % 5 == 0) ? 1.f : 0.f;
    }
    for (int p = 0; p < np; ++p) {
        const int k = g.pnp_kf[(size_t)p];
        const int i = g.kf_local[(size_t)k];
        SeqState& st = seq_[(size_t)g.seqs[(size_t)i]];
        const svo_pnp_result& r = g.pnp_res[(size_t)p];
        // solvePnPRansac leaves rvec/tvec as they were when it fails (the reference does not look
        // at its return value); either way they are the next keyframe's guess
        std::memcpy(st.rt, r.rvec, 3 * sizeof(double));
        std::memcpy(st.rt + 3, r.tvec, 3 * sizeof(double));
        const int ni = r.ok ? r.n_inliers : 0;
        g.n_inl[(size_t)k] = ni;
        st.res[slot].n_inliers = ni;
        for (int j = 0; j < ni; ++j) {
            const int src = g.pnp_inl[(size_t)p * mf + j];
            g.inl[(size_t)k * mf + j] = g.tracked[(size_t)i * mf + src];
            g.ids[(size_t)k * mf + j] = g.tracked_ids[(size_t)i * mf + src];
        }
        float* pose = &g.poses[(size_t)k * 16];           // hmat = [R^T | -R^T t]
        for (int a = 0; a < 3; ++a) {
            double acc = 0.0;
            for (int b = 0; b < 3; ++b) {
                pose[4 * a + b] = (float)r.rmat[3 * b + a];
                acc += r.rmat[3 * b + a] * r.tvec[b];
            }
            pose[4 * a + 3] = (float)(-acc);
        }
    }
    int rc = svo_keyframe(e, nk, g.kf_local.data(), g.inl.data(), g.n_inl.data(), g.poses.data(),
                          opt_.focal, opt_.cx, opt_.cy, opt_.baseline, g.new_feats.data(),
                          g.n_newv.data(), g.valid_2d.data(), g.pts_3d.data(), g.valid_idx.data(),
                          g.disp16.data(), g.n_valid.data());
    if (rc != SVO_OK) throw std::runtime_error(engine_error(e, "svo_keyframe", rc));
    // ---- add_keyframe's ids (src/bundle_adjuster.cpp:98-116) and FeatureTracker::init (:148-162)
    for (int k = 0; k < nk; ++k) {
        const int i = g.kf_local[(size_t)k];
        const int s = g.seqs[(size_t)i];
        SeqState& st = seq_[(size_t)s];
        SeqResult& rs = st.res[slot];
        const int n_inl = g.n_inl[(size_t)k];
        const int nv = std::min(g.n_valid[(size_t)k], mf - n_inl);
        svo_float2* f = &g.feats[(size_t)k * mf];
        uint64_t* id = &g.ids[(size_t)k * mf];
        std::memcpy(f, &g.inl[(size_t)k * mf], (size_t)n_inl * sizeof(svo_float2));
        std::memcpy(f + n_inl, &g.valid_2d[(size_t)k * mc], (size_t)nv * sizeof(svo_float2));
        for (int j = 0; j < nv; ++j) {
            id[n_inl + j] = st.next_id++;
            st.world.push_back(g.pts_3d[(size_t)k * mc + j]);
        }
        g.counts[(size_t)k] = n_inl + nv;
        std::memcpy(st.hmat, &g.poses[(size_t)k * 16], sizeof st.hmat);
        st.have_kf = true;
        rs.n_new = nv;
        rs.n_new_features = g.n_newv[(size_t)k];
        rs.n_valid = g.n_valid[(size_t)k];
        st.n_tracker = rs.n_tracker = n_inl + nv;
        std::memcpy(&tracker_pts_[slot][(size_t)s * mf], f, (size_t)(n_inl + nv) * sizeof(svo_float2));
        std::memcpy(&tracker_ids_[slot][(size_t)s * mf], id, (size_t)(n_inl + nv) * sizeof(uint64_t));
        std::memcpy(&new_2d_[slot][(size_t)s * mc], &g.valid_2d[(size_t)k * mc], (size_t)nv * sizeof(svo_float2));
        std::memcpy(&new_3d_[slot][(size_t)s * mc], &g.pts_3d[(size_t)k * mc], (size_t)nv * sizeof(svo_float3));
        std::memcpy(&new_feats_[slot][(size_t)s * mc], &g.new_feats[(size_t)k * mc], (size_t)rs.n_new_features * sizeof(svo_float2));
        std::memcpy(&disp16_[slot][(size_t)s * mc], &g.disp16[(size_t)k * mc], (size_t)rs.n_new_features * sizeof(int16_t));
        std::memcpy(&valid_idx_[slot][(size_t)s * mc], &g.valid_idx[(size_t)k * mc], (size_t)rs.n_valid * sizeof(int32_t));
        sum.keyframes += 1;
        sum.new_points += nv;
    }
    rc = svo_tracker_set(e, nk, g.kf_local.data(), g.feats.data(), g.ids.data(), g.counts.data());
    if (rc != SVO_OK) throw std::runtime_error(engine_error(e, "svo_tracker_set", rc));
}

void BatchFrontEnd::set_closed_loop(bool on, int iterations, float reproj_error, double confidence) {
    if (n_submitted_ != 0) throw std::logic_error("BatchFrontEnd::set_closed_loop: call before the first step");
    closed_loop_ = on;
    pnp_iterations_ = iterations;
    pnp_reproj_ = reproj_error;
    pnp_confidence_ = confidence;
}

void BatchFrontEnd::submit(const uint8_t* const* left, const uint8_t* const* right, int pitch,
                           bool on_device, const float* poses, const uint8_t* const* next_left,
                           const uint8_t* const* next_right) {
    if (!left) throw std::invalid_argument("BatchFrontEnd::submit: left is null");
    if (in_flight() >= kPipeDepth)
        throw std::logic_error("BatchFrontEnd::submit: pipeline full, call wait() first");
    const size_t n = (size_t)opt_.n_seq;
    Job& job = jobs_[n_submitted_ % kPipeDepth];
    job.left.assign(left, left + n);
    job.has_right = right != nullptr;
    if (right) job.right.assign(right, right + n);
    job.has_next = next_left != nullptr;
    if (next_left) job.next_left.assign(next_left, next_left + n);
    job.has_next_right = next_left != nullptr && next_right != nullptr;
    if (job.has_next_right) job.next_right.assign(next_right, next_right + n);
    job.has_poses = poses != nullptr;
    if (poses) job.poses.assign(poses, poses + n * 16);
    job.pitch = pitch;
    job.on_device = on_device;
    {
        std::lock_guard<std::mutex> lk(mu_);
        job.pending = (int)groups_.size();
        ++n_submitted_;
    }
    cv_start_.notify_all();
}

BatchSummary BatchFrontEnd::wait() {
    if (in_flight() == 0) throw std::logic_error("BatchFrontEnd::wait: nothing submitted");
    const int slot = (int)(n_waited_ % kPipeDepth);
    Job& job = jobs_[slot];
    {
        std::unique_lock<std::mutex> lk(mu_);
        cv_done_.wait(lk, [&] { return job.pending == 0; });
    }
    ++n_waited_;
    cur_slot_ = slot;
    BatchSummary total;
    for (size_t g = 0; g < groups_.size(); ++g) {
        if (!job.error[g].empty()) throw std::runtime_error(job.error[g]);
        const BatchSummary& sm = job.sum[g];
        total.frames += sm.frames; total.keyframes += sm.keyframes; total.corners += sm.corners;
        total.tracked += sm.tracked; total.new_points += sm.new_points;
        total.right_images += sm.right_images;
    }
    return total;
}

BatchSummary BatchFrontEnd::step(const uint8_t* const* left, const uint8_t* const* right, int pitch,
                                 bool on_device, const float* poses,
                                 const uint8_t* const* next_left,
                                 const uint8_t* const* next_right) {
    while (in_flight() > 0) wait();      // a synchronous step never overtakes queued ones
    submit(left, right, pitch, on_device, poses, next_left, next_right);
    return wait();
}

const svo_float2* BatchFrontEnd::corners(int s) const { return &corners_[cur_slot_][(size_t)s * opt_.params.max_corners]; }
const svo_float2* BatchFrontEnd::tracker_points(int s) const { return &tracker_pts_[cur_slot_][(size_t)s * opt_.params.max_features]; }
const uint64_t* BatchFrontEnd::tracker_ids(int s) const { return &tracker_ids_[cur_slot_][(size_t)s * opt_.params.max_features]; }
const svo_float2* BatchFrontEnd::new_2d(int s) const { return &new_2d_[cur_slot_][(size_t)s * opt_.params.max_corners]; }
const svo_float3* BatchFrontEnd::new_3d(int s) const { return &new_3d_[cur_slot_][(size_t)s * opt_.params.max_corners]; }
const svo_float2* BatchFrontEnd::new_features(int s) const { return &new_feats_[cur_slot_][(size_t)s * opt_.params.max_corners]; }
const int16_t* BatchFrontEnd::disp16(int s) const { return &disp16_[cur_slot_][(size_t)s * opt_.params.max_corners]; }
const int32_t* BatchFrontEnd::valid_idx(int s) const { return &valid_idx_[cur_slot_][(size_t)s * opt_.params.max_corners]; }

void BatchFrontEnd::set_blocking_sync(bool on) {
    for (auto& g : groups_) svo_set_blocking_sync(g->e, on ? 1 : 0);
}

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
