"""CPU: the image-message side of the reference's node without ROS (src/vo_node.cpp:38-76,
139-143; include/stereo_vo/image_msg.hpp): sensor_msgs/Image -> mono8 the way
cv_bridge::toCvCopy(msg, MONO8) does it (pinned against cv2: cvtColor for the colour encodings,
convertTo(255 / 65535) for the 16-bit ones), handle_images' drop rule and queue, the drain loop."""
import numpy as np
import cv2
import pytest


@pytest.fixture(scope="module")
def H():
    from stereo_vo_b200 import build, hostapi
    if build.host_needs_build():
        build.build_host_lib()
    hostapi.load_library()
    return hostapi


def _narrow(img16):
    """cv_bridge's 16 -> 8 bit step: convertTo(CV_8U, 255. / 65535.)"""
    return cv2.convertScaleAbs(img16, alpha=255.0 / 65535.0)


def test_mono8_is_a_row_by_row_copy(H):
    rng = np.random.default_rng(0)
    h, w, step = 37, 101, 128
    buf = rng.integers(0, 256, (h, step), dtype=np.uint8)
    out = H.to_mono8(H.image_msg(buf, "mono8", width=w))
    assert np.array_equal(out, buf[:, :w])
    out = H.to_mono8(H.image_msg(buf, "mono8", width=w), pitch=160)     # pitched destination
    assert np.array_equal(out, buf[:, :w])


@pytest.mark.parametrize("enc,code", [("rgb8", cv2.COLOR_RGB2GRAY), ("bgr8", cv2.COLOR_BGR2GRAY),
                                      ("rgba8", cv2.COLOR_RGBA2GRAY), ("bgra8", cv2.COLOR_BGRA2GRAY)])
def test_colour8_equals_cvtcolor(H, enc, code):
    rng = np.random.default_rng(1)
    ch = 4 if "a" in enc else 3
    img = rng.integers(0, 256, (53, 131, ch), dtype=np.uint8)
    img[0, :4] = [[255] * ch, [0] * ch, [255, 0, 0, 7][:ch], [0, 0, 255, 9][:ch]]
    ref = cv2.cvtColor(img, code)
    assert np.array_equal(H.to_mono8(H.image_msg(img, enc)), ref)
    # rows padded beyond width * channels
    padded = np.zeros((53, 131 * ch + 13), np.uint8)
    padded[:, :131 * ch] = img.reshape(53, -1)
    assert np.array_equal(H.to_mono8(H.image_msg(padded, enc, width=131)), ref)


def test_mono16_every_value_both_byte_orders(H):
    allv = np.arange(65536, dtype=np.uint16).reshape(256, 256)
    ref = _narrow(allv)
    little = allv.astype("<u2").view(np.uint8).reshape(256, 512)
    big = allv.astype(">u2").view(np.uint8).reshape(256, 512)
    assert np.array_equal(H.to_mono8(H.image_msg(little, "mono16", width=256)), ref)
    assert np.array_equal(H.to_mono8(H.image_msg(big, "mono16", width=256, big_endian=True)), ref)
    assert not np.array_equal(H.to_mono8(H.image_msg(big, "mono16", width=256)), ref)


@pytest.mark.parametrize("enc,code", [("rgb16", cv2.COLOR_RGB2GRAY), ("bgr16", cv2.COLOR_BGR2GRAY),
                                      ("rgba16", cv2.COLOR_RGBA2GRAY), ("bgra16", cv2.COLOR_BGRA2GRAY)])
def test_colour16_is_cvtcolor_then_narrowing(H, enc, code):
    rng = np.random.default_rng(2)
    ch = 4 if "a" in enc else 3
    img = rng.integers(0, 65536, (41, 67, ch), dtype=np.uint16)
    img[0, :2] = [[65535] * ch, [0] * ch]
    ref = _narrow(cv2.cvtColor(img, code))
    raw = img.astype("<u2").view(np.uint8).reshape(41, -1)
    assert np.array_equal(H.to_mono8(H.image_msg(raw, enc, width=67)), ref)
    raw_be = img.astype(">u2").view(np.uint8).reshape(41, -1)
    assert np.array_equal(H.to_mono8(H.image_msg(raw_be, enc, width=67, big_endian=True)), ref)


def test_messages_cv_bridge_would_refuse(H):
    from stereo_vo_b200.capi import SvoError
    buf = np.zeros((8, 64), np.uint8)
    for enc in ("8UC1", "32FC1", "bayer_rggb8", "yuv422", "", "MONO8"):
        with pytest.raises(SvoError, match="cannot convert encoding"):
            H.to_mono8(H.image_msg(buf, enc, width=8))
    with pytest.raises(SvoError, match="step is smaller"):
        H.to_mono8(H.image_msg(buf, "rgb8", width=30))              # needs 90 bytes per row
    with pytest.raises(SvoError, match="shorter than step"):
        m = H.image_msg(buf, "mono8", width=8)
        m.size = 8 * 64 - 1
        H.to_mono8(m)
    with pytest.raises(SvoError, match="empty image"):
        H.to_mono8(H.image_msg(buf, "mono8", width=0), pitch=8)
    with pytest.raises(SvoError, match="pitch < width"):
        H.to_mono8(H.image_msg(buf, "mono8", width=32), pitch=16)
    with pytest.raises(SvoError, match="null"):
        m = H.image_msg(buf, "mono8", width=8)
        m.data = None
        H.to_mono8(m)


def _pair(H, t, value, enc="mono8"):
    l = np.full((6, 10), value, np.uint8)
    r = np.full((6, 10), value + 1, np.uint8)
    if enc == "bgr8":
        r = np.repeat(r[:, :, None], 3, axis=2)
    return H.image_msg(l, "mono8", stamp=t), H.image_msg(r, enc, stamp=t + 0.001)


def test_handle_images_drop_rule_and_queue_order(H):
    """src/vo_node.cpp:64-73: a pair is dropped when the queue is not empty and its LEFT stamp is
    less than drop_time after the last QUEUED pair's; dropped pairs are not even converted."""
    q = H.StereoQueue(0.05)
    try:
        took = []
        for i, t in enumerate([10.00, 10.02, 10.05, 10.07, 10.0999, 10.11, 9.0, 10.30]):
            took.append(q.push(*_pair(H, t, 10 * i, "bgr8" if i % 2 else "mono8")))
        expect = []
        back = None
        for i, t in enumerate([10.00, 10.02, 10.05, 10.07, 10.0999, 10.11, 9.0, 10.30]):
            keep = back is None or not (t - back < 0.05)
            expect.append(keep)
            if keep:
                back = t
        assert took == expect and sum(took) == 4
        assert len(q) == 4
        # a dropped pair is never converted: a bad encoding goes unnoticed (return at :66)
        bad = H.image_msg(np.zeros((6, 10), np.uint8), "32FC1", stamp=10.31)
        assert q.push(bad, bad) is False
        # ... and a kept one with a bad encoding raises and queues nothing
        from stereo_vo_b200.capi import SvoError
        bad = H.image_msg(np.zeros((6, 10), np.uint8), "32FC1", stamp=11.0)
        with pytest.raises(SvoError, match="cannot convert encoding"):
            q.push(bad, bad)
        assert len(q) == 4
        seen = []
        n = q.drain(lambda l, r, t: seen.append((t, int(l[0, 0]), int(r[0, 0]), l.shape, r.shape)))
        assert n == 4 and len(q) == 0
        assert [s[0] for s in seen] == [10.00, 10.05, 10.11, 10.30]        # oldest first
        assert [s[1] for s in seen] == [0, 20, 50, 70] and [s[2] for s in seen] == [1, 21, 51, 71]
        assert all(s[3] == (6, 10) and s[4] == (6, 10) for s in seen)
        assert q.drain(lambda l, r, t: 0) == 0                             # empty queue: no calls
        # an empty queue takes any stamp (the rule needs a previous pair)
        assert q.push(*_pair(H, 1.0, 5)) is True
    finally:
        q.close()


def test_drain_stops_at_a_failing_consumer_and_keeps_the_pair(H):
    from stereo_vo_b200.capi import SvoError
    q = H.StereoQueue(0.0)                    # drop_time 0: nothing is dropped going forwards
    try:
        for i in range(5):
            assert q.push(*_pair(H, float(i), i))
        seen = []

        def consumer(l, r, t):
            if t == 2.0:
                return 7
            seen.append(t)

        with pytest.raises(SvoError, match="consumer returned 7"):
            q.drain(consumer)
        assert seen == [0.0, 1.0] and q.last_drained == 2 and len(q) == 3
        rest = []
        assert q.drain(lambda l, r, t: rest.append(t)) == 3 and rest == [2.0, 3.0, 4.0]
    finally:
        q.close()


def test_queued_images_are_copies(H):
    """cv_bridge::toCvCopy: the queue owns its pixels; the message buffer may be reused."""
    q = H.StereoQueue(0.05)
    try:
        l = np.full((4, 8), 9, np.uint8)
        r = np.full((4, 8), 3, np.uint8)
        q.push(H.image_msg(l, "mono8", stamp=1.0), H.image_msg(r, "mono8", stamp=1.0))
        l[:] = 200
        r[:] = 201
        got = []
        q.drain(lambda a, b, t: got.append((a.copy(), b.copy())))
        assert (got[0][0] == 9).all() and (got[0][1] == 3).all()
    finally:
        q.close()
