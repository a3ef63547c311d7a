"""GPU parity: pyramidal LK forward/backward + filter/statistics (FeatureTracker) against the
cv2 oracle. Status, kept indices and ids bit-exact; sub-pixel flow bit-exact (north_star allows 0.01 px)
(BASELINE.json north_star tolerance)."""
import numpy as np
import pytest

from stereo_vo_b200 import synth
from tests import util

pytestmark = pytest.mark.gpu
FLOW_TOL = 0.0           # px: flow is bit-exact (window sums in cv2's own f32 order)
PAR_TOL = 1e-4           # av_parallax: an f32 mean of f32 norms


@pytest.fixture(scope="module")
def eng(svo_lib):
    from stereo_vo_b200.capi import Engine
    cache = {}

    def get(w, h, **kw):
        key = (w, h, tuple(sorted(kw.items())))
        if key not in cache:
            cache[key] = Engine(w, h, n_seq=1, **kw)
        return cache[key]
    yield get
    for e in cache.values():
        e.close()


def _track_once(e, oracle, prev, nxt, pts, ids):
    trk = oracle.FeatureTracker()
    trk.init(prev, pts, ids)
    av_ref, lost_ref = trk.track_features(nxt, True)
    e.tracker_init(prev, pts, np.asarray(ids, np.uint64))
    av, lost = e.tracker_track(nxt, True)
    dbg = e.tracker_debug()
    taps = trk.taps
    assert np.array_equal(dbg["status1"], taps["status1"]), "forward status"
    ok1 = taps["status1"] > 0
    assert np.abs(dbg["tracked"][ok1] - taps["tracked"][ok1]).max() <= FLOW_TOL
    # backward pass starts from OUR forward result; compare where both forward results agree
    assert np.array_equal(dbg["status2"][ok1], taps["status2"][ok1]), "backward status"
    ok2 = ok1 & (taps["status2"] > 0)
    assert np.abs(dbg["reverse"][ok2] - taps["reverse"][ok2]).max() <= 2 * FLOW_TOL
    feats, fids = e.tracker_get()
    ref_feats, ref_ids = trk.get_tracked_features()
    assert list(fids) == list(ref_ids), "kept track ids"
    assert np.abs(feats - ref_feats).max() <= FLOW_TOL
    assert np.array_equal(dbg["kept_idx"][:len(ref_ids)], taps["kept_idx"])
    assert abs(float(av) - float(av_ref)) <= PAR_TOL
    assert lost == lost_ref
    return trk


@pytest.mark.parametrize("seed,t", [(0, 0), (1, 7), (2, 20)])
def test_track_kitti_pair(eng, oracle, seed, t):
    fr = util.frames(seed, 2, start=t)
    prev, nxt = fr[0][0], fr[1][0]
    pts = oracle.detect(prev)
    e = eng(1241, 376)
    _track_once(e, oracle, prev, nxt, pts, list(range(100, 100 + len(pts))))


def test_track_d435i_pair(eng, oracle):
    fr = synth.sequence(3, 2, synth.D435I, synth.D435I_MOTION)
    prev, nxt = fr[0][0], fr[1][0]
    pts = oracle.detect(prev)
    e = eng(640, 480)
    _track_once(e, oracle, prev, nxt, pts, list(range(len(pts))))


def test_track_border_points_and_lost(eng, oracle):
    """points at / beyond the image edge, a large shift that loses tracks, duplicate points"""
    fr = util.frames(4, 1)
    prev = fr[0][0]
    nxt = np.roll(np.roll(prev, -14, axis=1), -9, axis=0)
    pts = oracle.detect(prev)[:60]
    edge = np.array([[0, 0], [1240, 375], [3.5, 200.25], [1238.75, 10.5], [620, 0.5],
                     [620, 374.9], [0.2, 0.3], [1240.0, 0.0], [5, 370], [1230, 5]], np.float32)
    pts = np.concatenate([pts, edge, pts[:3]], 0)
    e = eng(1241, 376)
    _track_once(e, oracle, prev, nxt, pts, list(range(len(pts))))


def test_track_flat_image_all_lost(eng, oracle):
    prev = np.full((376, 1241), 120, np.uint8)
    nxt = prev.copy()
    pts = np.array([[100, 100], [600, 200], [1200, 300]], np.float32)
    trk = oracle.FeatureTracker()
    trk.init(prev, pts, [0, 1, 2])
    av_ref, lost_ref = trk.track_features(nxt, True)
    e = eng(1241, 376)
    e.tracker_init(prev, pts, np.arange(3, dtype=np.uint64))
    av, lost = e.tracker_track(nxt, True)
    assert len(trk.feature_ids) == 0 and e.tracker_get()[0].shape[0] == 0
    assert av == av_ref and lost == lost_ref == np.float32(1.0)


def test_track_multi_frame_chain(eng, oracle):
    """four consecutive frames: the tracker state (kept set, ids, initial positions) carries over"""
    fr = util.frames(5, 5)
    pts = oracle.detect(fr[0][0])
    ids = list(range(len(pts)))
    trk = oracle.FeatureTracker()
    trk.init(fr[0][0], pts, ids)
    e = eng(1241, 376)
    e.tracker_init(fr[0][0], pts, np.asarray(ids, np.uint64))
    for t in range(1, 5):
        av_ref, lost_ref = trk.track_features(fr[t][0], True)
        av, lost = e.tracker_track(fr[t][0], True)
        feats, fids = e.tracker_get()
        assert list(fids) == list(trk.feature_ids), f"frame {t}"
        assert np.abs(feats - trk.feature_set).max() <= FLOW_TOL
        assert abs(float(av) - float(av_ref)) <= PAR_TOL
        assert lost == lost_ref



def test_backward_to_forward_handover_is_used_and_changes_nothing(svo_lib, oracle):
    """lk_track hands the window state of the backward pass to the next frame's forward pass
    (svo_lk_handover). Same frames through an engine with the hand-over and one without: every
    LK output bit for bit equal; the hand-over engine must actually hit (every kept feature of
    every frame after the first tracked one), and newly installed feature sets must miss."""
    from stereo_vo_b200.capi import Engine
    fr = util.frames(6, 8)
    pts = oracle.detect(fr[0][0])
    ids = np.arange(len(pts), dtype=np.uint64)
    on, off = Engine(1241, 376, n_seq=1), Engine(1241, 376, n_seq=1)
    try:
        on.lk_handover(enable=1, count=1)
        off.lk_handover(enable=0, count=1)
        for e in (on, off):
            e.tracker_init(fr[0][0], pts, ids)
        kept_before = []
        for t in range(1, 8):
            kept_before.append(on.tracker_get()[0].shape[0])
            ra, rb = on.tracker_track(fr[t][0], True), off.tracker_track(fr[t][0], True)
            assert ra == rb, f"frame {t}: statistics"
            da, db = on.tracker_debug(), off.tracker_debug()
            for key in ("tracked", "status1", "reverse", "status2", "kept_idx"):
                assert np.array_equal(da[key], db[key]), f"frame {t}: {key}"
            fa, fb = on.tracker_get(), off.tracker_get()
            assert np.array_equal(fa[0], fb[0]) and list(fa[1]) == list(fb[1]), f"frame {t}"
            if t == 4:      # the host installs a feature set: look-ups of the next frame must miss
                for e in (on, off):
                    e.tracker_init(fr[t][0], fa[0][::-1].copy(), np.asarray(fa[1][::-1], np.uint64))
        lookups, hits = on.lk_handover()
        # frame 1 (no backward pass yet) and frame 5 (set installed by the host, reversed order:
        # the index map no longer leads to the entries) recompute; frames 2-4, 6, 7 take it
        assert lookups == sum(kept_before)
        expect = sum(kept_before[i] for i in (1, 2, 3, 5, 6))
        assert hits >= expect - 2 and hits <= expect + kept_before[4], (lookups, hits, kept_before)
        assert off.lk_handover() == (0, 0)
    finally:
        on.close()
        off.close()


def test_handover_survives_a_feature_set_installed_by_the_host(svo_lib, oracle, monkeypatch):
    """svo_tracker_set (the facade's FeatureTracker::init on the resident frame, the closed loop
    after every keyframe) re-points the hand-over index map at the entries of the features the
    host kept (lk_rekey_kernel): a reordered subset of the tracked points must hit on the next
    frame, points no entry exists for must miss, and every LK output stays bit for bit what an
    engine without the hand-over computes."""
    from stereo_vo_b200.capi import Engine
    fr = util.frames(7, 6)
    pts = oracle.detect(fr[0][0])
    ids = np.arange(len(pts), dtype=np.uint64)
    n_fresh = 40
    monkeypatch.setenv("SVO_LK_REKEY", "1")          # read by svo_create
    on, off = Engine(1241, 376, n_seq=1), Engine(1241, 376, n_seq=1)
    try:
        on.lk_handover(enable=1, count=1)
        off.lk_handover(enable=0, count=1)
        for e in (on, off):
            e.tracker_init(fr[0][0], pts, ids)
        looked, expect = 0, 0
        can_hit = 0                      # frame 1: no backward pass has run yet
        for t in range(1, 6):
            looked += on.tracker_get()[0].shape[0]
            expect += can_hit
            ra, rb = on.tracker_track(fr[t][0], True), off.tracker_track(fr[t][0], True)
            assert ra == rb, f"frame {t}: statistics"
            da, db = on.tracker_debug(), off.tracker_debug()
            for key in ("tracked", "status1", "reverse", "status2"):
                assert np.array_equal(da[key], db[key]), f"frame {t}: {key}"
            fa, fb = on.tracker_get(), off.tracker_get()
            assert np.array_equal(fa[0], fb[0]) and list(fa[1]) == list(fb[1]), f"frame {t}"
            can_hit = fa[0].shape[0]     # the filter's own index map
            if t in (2, 3):
                # the host keeps every second point, in reverse order; after frame 3 it also
                # adds points that were never tracked
                keep = fa[0][::2][::-1].copy()
                keep_ids = np.asarray(fa[1][::2][::-1], np.uint64)
                can_hit = keep.shape[0]
                assert can_hit > 20
                if t == 3:
                    fresh = oracle.detect(fr[t][0])[:n_fresh] + np.float32(0.25)
                    assert fresh.shape[0] == n_fresh
                    keep = np.concatenate([keep, fresh], 0)
                    keep_ids = np.concatenate(
                        [keep_ids, 10_000 + np.arange(n_fresh, dtype=np.uint64)])
                for e in (on, off):
                    e.tracker_set([0], [keep], [keep_ids])
        lookups, hits = on.lk_handover()
        assert lookups == looked
        # (a point whose window left the bordered image on some level has no complete entry)
        assert expect - 4 <= hits <= expect, (lookups, hits, expect)
        assert off.lk_handover() == (0, 0)
    finally:
        on.close()
        off.close()
