"""GPU debug overlay (svo_draw_track / FeatureTracker::draw_track) against
the reference's own calls: cv2.cvtColor(GRAY2RGB) + cv2.arrowedLine(..., (0, 255, 0), 4)
(src/feature_tracker.cpp:74-83). Untouched pixels must be the grey image exactly; the painted
mask must be OpenCV's up to its outline (OpenCV rasterises the thick line as a fixed-point
polygon plus round caps, the kernel as a distance test)."""
import cv2
import numpy as np
import pytest

from stereo_vo_b200 import synth
from tests import util

pytestmark = pytest.mark.gpu


def _reference_overlay(gray, a, b):
    img = cv2.cvtColor(gray, cv2.COLOR_GRAY2RGB)
    for p, q in zip(a, b):
        cv2.arrowedLine(img, (int(np.rint(p[0])), int(np.rint(p[1]))), (int(np.rint(q[0])), int(np.rint(q[1]))),
                        (0, 255, 0), 4)
    return img


def _masks_agree(got, ref, gray):
    green = np.array([0, 255, 0], np.uint8)
    mg, mr = np.all(got == green, axis=2), np.all(ref == green, axis=2)
    rgb = np.dstack([gray] * 3)
    assert np.array_equal(got[~mg], rgb[~mg]), "unpainted pixels must be the grey image"
    inter, union = (mg & mr).sum(), (mg | mr).sum()
    k = np.ones((3, 3), np.uint8)
    grow_g = cv2.dilate(mg.astype(np.uint8), k) > 0
    grow_r = cv2.dilate(mr.astype(np.uint8), k) > 0
    assert not (mr & ~grow_g).any(), "OpenCV paints pixels more than 1 px away from the kernel's"
    assert not (mg & ~grow_r).any(), "the kernel paints pixels more than 1 px away from OpenCV's"
    return inter / max(union, 1)


def test_arrows_match_cv2(svo_lib):
    from stereo_vo_b200.capi import Engine
    cam = synth.KITTI00
    gray = util.frames(4, 1)[0][0]
    rng = np.random.default_rng(8)
    n = 300
    a = np.stack([rng.uniform(0, cam.width, n), rng.uniform(0, cam.height, n)], 1).astype(np.float32)
    b = (a + rng.normal(scale=18.0, size=(n, 2))).astype(np.float32)
    b[:10] = a[:10]                                         # zero-length tracks
    b[10:20] += 400                                         # arrows that leave the image
    e = Engine(cam.width, cam.height, n_seq=1)
    got = e.draw_track(gray, a, b)
    iou = _masks_agree(got, _reference_overlay(gray, a, b), gray)
    print(f"\n[overlay] {n} arrows: painted-mask IoU with cv2.arrowedLine {iou:.3f}")
    assert iou >= 0.85
    assert np.array_equal(e.draw_track(gray, a[:0], b[:0]), cv2.cvtColor(gray, cv2.COLOR_GRAY2RGB))
    e.close()


def test_facade_drawing_after_keyframes(svo_lib):
    """ImageProcessor::process calls draw_track on keyframes (src/image_processor.cpp:146);
    get_drawing returns the RGB overlay vo_node publishes (vo_node.cpp:188)."""
    from stereo_vo_b200 import build, hostapi
    if build.host_needs_build():
        build.build_host_lib()
    cam = synth.KITTI00
    f, cx, cy, b = cam.f32()
    vo = hostapi.VoSequence(f, cx, cy, b)
    fr = util.frames(0, 8)
    first_kf = None
    for t, (l, r) in enumerate(fr):
        rec = vo.process(l, r, 0.1 * t)
        if rec["decision"] == "keyframe" and first_kf is None:
            first_kf = t
            img = vo.get_drawing()
            assert img.shape == (cam.height, cam.width, 3)
            green = np.all(img == np.array([0, 255, 0], np.uint8), axis=2)
            assert green.sum() > 50 * 20                    # hundreds of arrows, ~ 20 px long
            # the background is the frame the tracks started from (the previous keyframe)
            assert np.array_equal(img[~green], np.dstack([fr[0][0]] * 3)[~green])
    assert first_kf is not None
    vo.close()
