"""svo_dedupe_static (include/svo_c.h): the allocation-free variant of svo_dedupe, whose scratch is
a static device array inside libsvo.so. No counterpart in the reference.
Called through the C ABI and compared with svo_dedupe on the same inputs -
kept points and kept indices bit for bit, as the other dedupe tests demand - on two calls in a
row (the second sees the scratch the first one left), then at the stated maximum, then past it.

Named test_zz_* so that pytest collects it last.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case(rng, n_det, n_trk, w=1241, h=376):
    """integer corner positions (what the detector yields) against sub-pixel tracked points,
    a share of them planted right at the min-distance border of a detection"""
    det = np.stack([rng.integers(0, w, n_det), rng.integers(0, h, n_det)], 1).astype(np.float32)
    trk = np.stack([rng.uniform(0, w, n_trk), rng.uniform(0, h, n_trk)], 1).astype(np.float32)
    k = min(n_det, n_trk) // 3
    if k:
        ang = rng.uniform(0, 2 * np.pi, k)
        rad = rng.choice([29.0, 29.999, 30.0, 30.001, 31.0, 5.0], k)
        trk[:k] = det[rng.choice(n_det, k, replace=False)] + (np.stack([np.cos(ang), np.sin(ang)], 1) * rad[:, None]).astype(np.float32)
    return det, trk


class _Pinned:
    """caller-owned buffers the device can dereference (svo_host_alloc): allocated ONCE, before
    any call of the entry point under test"""

    def __init__(self, cap_det, cap_trk):
        from stereo_vo_b200 import capi
        self.det = capi.pinned_empty((cap_det, 2), np.float32)
        self.trk = capi.pinned_empty((cap_trk, 2), np.float32)
        self.out = capi.pinned_empty((cap_det, 2), np.float32)
        self.keep = capi.pinned_empty((cap_det,), np.int32)
        self.n = capi.pinned_empty((1,), np.int32)

    def run(self, det, trk, md):
        from stereo_vo_b200 import capi
        nd, nt = len(det), len(trk)
        self.det[:nd] = det
        self.trk[:nt] = trk
        self.out[:] = -7.0
        self.keep[:] = -7
        self.n[0] = -7
        capi.dedupe_static(self.det, nd, self.trk, nt, md, self.out, self.keep, self.n)
        m = int(self.n[0])
        assert 0 <= m <= nd
        return self.out[:m].copy(), self.keep[:m].copy()


def test_dedupe_static_matches_svo_dedupe_twice_in_a_row(svo_lib):
    from stereo_vo_b200.capi import Engine
    rng = np.random.default_rng(2024)
    e = Engine(1241, 376, n_seq=1, max_corners=4096, max_features=4096)
    bufs = _Pinned(4096, 4096)
    # two calls in a row with different sizes: the second must not see the first one's flags
    for call, (nd, nt, md) in enumerate([(3000, 60, 30.0), (1700, 2500, 30.0), (1, 1, 30.0),
                                         (900, 0, 30.0), (0, 50, 30.0), (2048, 333, 7.5)]):
        det, trk = _case(rng, max(nd, 1), max(nt, 1))
        det, trk = det[:nd], trk[:nt]
        ref_xy, ref_keep = e.dedupe(det, trk, md)
        got_xy, got_keep = bufs.run(det, trk, md)
        assert np.array_equal(got_keep, ref_keep), f"call {call}: kept indices differ"
        assert np.array_equal(got_xy.view(np.uint32), ref_xy.view(np.uint32)), f"call {call}: kept points differ"
        if nd >= 100 and nt >= 50 and md >= 10:
            assert len(ref_keep) < nd, f"call {call}: nothing was dropped"
        if call == 0:
            assert 200 < len(ref_keep), "call 0: nothing was kept"
    # the same input again: same answer (nothing carried over in the scratch)
    a = bufs.run(det, trk, md)
    b = bufs.run(det, trk, md)
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[0], b[0])
    e.close()


def test_dedupe_static_at_the_limit_and_beyond(svo_lib):
    """SVO_DEDUPE_STATIC_MAX_DET detections are served (checked against a numpy statement of the
    rule's threshold test in fp32); one more is refused with a message and nothing is written."""
    from stereo_vo_b200 import capi
    cap = capi.DEDUPE_STATIC_MAX_DET
    rng = np.random.default_rng(7)
    det, trk = _case(rng, cap, 300)
    bufs = _Pinned(cap + 1, 300)
    got_xy, got_keep = bufs.run(det, trk, 30.0)
    dx = det[:, None, 0] - trk[None, :, 0]
    dy = det[:, None, 1] - trk[None, :, 1]
    d2 = (dx * dx).astype(np.float32) + (dy * dy).astype(np.float32)
    near = (np.sqrt(d2.astype(np.float32)) < np.float32(30.0)).any(1)
    assert np.array_equal(got_keep, np.nonzero(~near)[0].astype(np.int32))
    assert np.array_equal(got_xy, det[~near])
    bufs.n[0] = -7
    with pytest.raises(capi.SvoError) as ei:
        capi.dedupe_static(bufs.det, cap + 1, bufs.trk, 300, 30.0, bufs.out, bufs.keep, bufs.n)
    assert ei.value.code == -3 and str(cap) in str(ei.value) and "exceeds" in str(ei.value)
    assert bufs.n[0] == -7
    # pageable host memory is refused, not dereferenced
    pageable = np.zeros((16, 2), np.float32)
    with pytest.raises(capi.SvoError) as ei:
        capi.dedupe_static(pageable, 16, bufs.trk, 300, 30.0, bufs.out, bufs.keep, bufs.n)
    assert ei.value.code == -1 and "pinned" in str(ei.value)
