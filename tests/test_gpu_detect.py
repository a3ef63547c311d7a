"""GPU parity: pyramid, Shi-Tomasi eig map, NMS candidates and goodFeaturesToTrack corners,
through the C ABI, against the cv2 oracle (bit-exact; integer / index work)."""
import cv2
import numpy as np
import pytest

from stereo_vo_b200 import synth
from tests import util

pytestmark = pytest.mark.gpu

SIZES = [(1241, 376), (640, 480), (1242, 375), (333, 211)]


@pytest.fixture(scope="module")
def engines(svo_lib):
    from stereo_vo_b200.capi import Engine
    cache = {}

    def get(w, h, **kw):
        key = (w, h, tuple(sorted(kw.items())))
        if key not in cache:
            cache[key] = Engine(w, h, n_seq=2, **kw)
        return cache[key]
    yield get
    for e in cache.values():
        e.close()


def _image(w, h, seed):
    if (w, h) == (1241, 376):
        return util.frames(seed, 1)[0][0]
    cam = util.small_camera(w, h)
    return synth.stereo_frame(seed, 3, cam)[0]


@pytest.mark.parametrize("w,h", SIZES)
def test_pyramid_levels_bit_exact(engines, oracle, w, h):
    e = engines(w, h)
    img = _image(w, h, 1)
    e.good_features_to_track(img)
    _, pyr = cv2.buildOpticalFlowPyramid(img, (21, 21), 3, withDerivatives=False)
    assert e.num_levels == len(pyr)
    ref = img
    for l in range(e.num_levels):
        got = e.debug_pyramid_level(l)
        if l > 0:
            ref = cv2.pyrDown(ref)
        assert got.shape == ref.shape
        assert np.array_equal(got, ref), f"level {l}"


@pytest.mark.parametrize("w,h", SIZES)
@pytest.mark.parametrize("seed", [0, 5])
def test_min_eigen_map_bit_exact(engines, oracle, w, h, seed):
    e = engines(w, h, keep_eig_map=1)
    img = _image(w, h, seed)
    e.good_features_to_track(img, seq=1)
    got = e.debug_min_eigen_map(seq=1)
    ref = cv2.cornerMinEigenVal(img, 3, ksize=3)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), \
        f"{(got != ref).sum()} pixels differ, max |d| = {np.abs(got - ref).max()}"


@pytest.mark.parametrize("w,h", SIZES)
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_good_features_bit_exact(engines, oracle, w, h, seed):
    e = engines(w, h)
    img = _image(w, h, seed)
    got = e.good_features_to_track(img)
    ref = oracle.detect(img)
    assert got.shape == ref.shape
    assert np.array_equal(got, ref)


def test_candidates_sorted_like_oracle(engines, oracle):
    from oracle import restate
    w, h = 1241, 376
    e = engines(w, h)
    img = _image(w, h, 3)
    e.good_features_to_track(img)
    got = e.debug_candidates()
    xs, ys, _ = restate.corner_candidates(cv2.cornerMinEigenVal(img, 3, ksize=3), 0.1)
    assert got.shape[0] == len(xs)
    assert np.array_equal(got[:, 0], xs) and np.array_equal(got[:, 1], ys)


def test_plateau_ties_and_flat_image(engines, oracle):
    w, h = 640, 480
    e = engines(w, h)
    flat = np.full((h, w), 93, np.uint8)
    assert e.good_features_to_track(flat).shape[0] == 0          # max == 0 -> no corners
    assert oracle.detect(flat).shape[0] == 0
    # blocky image: exact plateaus in the eig map (ties broken by address)
    rng = np.random.default_rng(7)
    blocks = rng.integers(0, 256, (h // 8, w // 8)).astype(np.uint8)
    img = np.kron(blocks, np.ones((8, 8), np.uint8))
    got = e.good_features_to_track(img)
    ref = oracle.detect(img)
    assert np.array_equal(got, ref)
    stress = util.stress_image(h, w, 11)
    assert np.array_equal(e.good_features_to_track(stress), oracle.detect(stress))


def _plateau_images(w, h):
    """images whose NMS survivors overflow the engine's candidate lists: period-3 patterns (the
    3x3 box sums see one whole period: the corner score is one plateau, EVERY pixel is a
    non-strict local maximum) and white noise (more than max_candidates = 16384 survivors)"""
    yy, xx = np.mgrid[0:h, 0:w]
    rng = np.random.default_rng(1)
    return {
        "dots3": np.where((xx % 3 == 0) & (yy % 3 == 0), 200, 20).astype(np.uint8),
        "tile3": np.array([[10, 200, 90], [160, 30, 220], [75, 240, 5]], np.uint8)[yy % 3, xx % 3],
        "noise": rng.integers(0, 256, (h, w)).astype(np.uint8),
    }


@pytest.mark.parametrize("w,h", [(1241, 376), (333, 211)])
def test_plateau_images_take_the_spill_path(engines, oracle, w, h):
    """cv::goodFeaturesToTrack has no candidate limit: where the engine's lists overflow, the
    detection is repeated on lists with room for every pixel and the corners are cv2's."""
    from oracle import restate
    e = engines(w, h)
    imgs = _plateau_images(w, h)
    normal = _image(w, h, 2)
    for name, img in imgs.items():
        ref = oracle.detect(img)
        before = e.detector_spills
        got = e.good_features_to_track(img)
        spilled = e.detector_spills - before
        assert np.array_equal(got, ref), f"{name}: corners differ ({len(got)} vs {len(ref)})"
        xs, ys, _ = restate.corner_candidates(cv2.cornerMinEigenVal(img, 3, ksize=3), 0.1)
        if len(xs) > 16384:
            assert spilled == 1, f"{name}: {len(xs)} survivors but no spill"
        cand = e.debug_candidates()
        assert cand.shape[0] == len(xs), name
        assert np.array_equal(cand[:, 0], xs) and np.array_equal(cand[:, 1], ys), f"{name}: candidate order"
        # through svo_step, next to an ordinary frame in the same call
        before = e.detector_spills
        res = e.step([0, 1], [normal, img], [normal, img])
        for i, im in enumerate((normal, img)):
            r = oracle.detect(im)
            n = res["n_corners"][i]
            assert n == r.shape[0] and np.array_equal(res["corners"][i, :n], r), f"{name}: step, sequence {i}"
        assert e.detector_spills - before == spilled
    # an ordinary frame afterwards does not spill
    before = e.detector_spills
    assert np.array_equal(e.good_features_to_track(normal), oracle.detect(normal))
    assert e.detector_spills == before
    assert e.detector_spills >= (2 if (w, h) == (1241, 376) else 1)


@pytest.mark.parametrize("maxc,q,md", [(50, 0.05, 10.0), (1000, 0.01, 5.0), (300, 0.3, 30.0),
                                       (200, 0.1, 0.0), (400, 0.02, 7.5)])
def test_other_detector_parameters(engines, oracle, maxc, q, md):
    w, h = 640, 480
    e = engines(w, h, max_corners=maxc, quality_level=q, min_distance=md)
    img = _image(w, h, 4)
    got = e.good_features_to_track(img)
    ref = oracle.detect(img, maxc, q, md)
    assert np.array_equal(got, ref)


def test_two_sequences_independent(engines, oracle):
    w, h = 1241, 376
    e = engines(w, h)
    a, b = _image(w, h, 8), _image(w, h, 9)
    res = e.step([0, 1], [a, b], [a, b])
    for i, img in enumerate((a, b)):
        ref = oracle.detect(img)
        n = res["n_corners"][i]
        assert n == ref.shape[0]
        assert np.array_equal(res["corners"][i, :n], ref)


_VARIANT_SCRIPT = r"""
import sys
import cv2
import numpy as np
sys.path.insert(0, {root!r})
from oracle import cv2_frontend as O
from stereo_vo_b200 import synth
from stereo_vo_b200.capi import Engine
from tests import util
O.canonical_mode()
for (w, h) in [(1241, 376), (640, 480), (333, 211)]:
    e = Engine(w, h, n_seq=2, keep_eig_map=1)
    for seed in (0, 3, 7):
        if (w, h) == (1241, 376):
            img = util.frames(seed, 1)[0][0]
        else:
            img = synth.stereo_frame(seed, 3, util.small_camera(w, h))[0]
        got = e.good_features_to_track(img, seq=1)
        ref = O.detect(img)
        assert got.shape == ref.shape and np.array_equal(got, ref), (w, h, seed, "corners")
        eig = e.debug_min_eigen_map(seq=1)
        ref_eig = cv2.cornerMinEigenVal(img, 3, ksize=3)
        assert np.array_equal(eig.view(np.uint32), ref_eig.view(np.uint32)), (w, h, seed, "eig")
    e.close()
print("VARIANT_OK")
"""


@pytest.mark.parametrize("env", [
    {"SVO_CORNER_FORCE_REPAIR": "1"},                                  # every segment re-walked
    {"SVO_CORNER_SEGMENTS": "1"},                                      # the sequential walk
    {"SVO_CORNER_SEGMENTS": "5", "SVO_CORNER_FORCE_REPAIR": "1"},
    {"SVO_CORNER_SEGMENTS": "16"},
    {"SVO_CORNER_CPT": "4", "SVO_CORNER_FORCE_REPAIR": "1"},
    {"SVO_CORNER_CPT": "1"},
    {"SVO_CORNER_MIN_BLOCKS": "20"},
], ids=lambda d: ",".join(f"{k[11:]}={v}" for k, v in d.items()))
def test_corner_score_variants(svo_lib, env):
    """corner_score_nms splits every strip into row segments that start from GUESSED sliding sums
    and repairs the ones whose guess was wrong; the variants (read from the environment once per
    process) force the repair pass, change the segmentation and the columns per thread - all
    must give the same bits as cv2."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    full_env = dict(os.environ, **env)
    r = subprocess.run([sys.executable, "-c", _VARIANT_SCRIPT.format(root=root)], env=full_env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "VARIANT_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def _hard_images(w, h):
    """Images built to make the sliding fp64 sums of cornerMinEigenVal round (values of very
    different magnitude in one column) and keep rounding residue alive across many rows: the
    cases in which the row segments of corner_score_nms start from a wrong guess."""
    rng = np.random.default_rng(2024)
    ys, xs = np.mgrid[0:h, 0:w]
    out = {}
    noise = rng.integers(0, 256, (h, w)).astype(np.uint8)
    half = noise.copy()
    half[:, w // 2:] = 131                                  # flat half beside full-range noise
    out["flat_half"] = half
    bands = noise.copy()
    for y0 in range(20, h, 70):                             # flat bands every 70 rows: residue
        bands[y0:y0 + 30, :] = 200                          # must survive into the next segment
    out["flat_bands"] = bands
    out["ramp_y"] = np.clip(ys * 255.0 / (h - 1), 0, 255).astype(np.uint8)       # Dx == 0: -0 xy sums
    out["ramp_x_sat"] = np.clip(xs * 400.0 / (w - 1), 0, 255).astype(np.uint8)  # saturated third
    out["checker"] = (((xs // 16) + (ys // 16)) % 2 * 255).astype(np.uint8)
    faint = np.full((h, w), 100, np.int32) + rng.integers(0, 2, (h, w))         # +-1 on a constant
    faint[h // 3:h // 3 + 9, w // 4:w // 4 + 9] = 255                             # one huge blob
    out["faint_plus_blob"] = faint.astype(np.uint8)
    mix = util.stress_image(h, w, 5).astype(np.int32)
    mix[:, ::97] = 0                                        # isolated extreme columns
    out["stress_columns"] = np.clip(mix, 0, 255).astype(np.uint8)
    return out


@pytest.mark.parametrize("w,h", [(1241, 376), (640, 480)])
def test_corner_segments_on_rounding_prone_images(engines, oracle, w, h):
    # full-range noise leaves more than the default 16384 NMS survivors: the tiled sort path
    e = engines(w, h, keep_eig_map=1, max_candidates=1 << 17)
    for name, img in _hard_images(w, h).items():
        got = e.good_features_to_track(img, seq=1)
        ref = oracle.detect(img)
        assert got.shape == ref.shape and np.array_equal(got, ref), name
        eig = e.debug_min_eigen_map(seq=1)
        ref_eig = cv2.cornerMinEigenVal(img, 3, ksize=3)
        assert np.array_equal(eig.view(np.uint32), ref_eig.view(np.uint32)), \
            f"{name}: {(eig.view(np.uint32) != ref_eig.view(np.uint32)).sum()} eig pixels differ"


def default_mode_agreement(detect_gpu, images):
    """Agreement of the GPU detector with cv2 in its DEFAULT (optimized) mode, the mode a stock
    OpenCV build runs in: its AVX2/AVX-512 cornerMinEigenVal contracts some multiply-adds into
    FMAs depending on the column, so a few eig values differ in the last
    bit from the canonical SSE path the GPU kernel reproduces. Returns (frames, frames with the
    identical corner SET, frames with the identical corner ORDER)."""
    was = cv2.useOptimized()
    same_set = same_order = 0
    try:
        for img in images:
            got = detect_gpu(img)
            cv2.setUseOptimized(True)
            ref = cv2.goodFeaturesToTrack(img, 300, 0.1, 30)
            cv2.setUseOptimized(False)
            ref = np.zeros((0, 2), np.float32) if ref is None else ref.reshape(-1, 2)
            s_got = {(float(x), float(y)) for x, y in got}
            s_ref = {(float(x), float(y)) for x, y in ref}
            same_set += s_got == s_ref
            same_order += got.shape == ref.shape and bool(np.array_equal(got, ref))
    finally:
        cv2.setUseOptimized(was)
    return len(images), same_set, same_order


def test_agreement_with_default_optimized_cv2_is_reported(engines, oracle):
    """Secondary number promised in SURVEY.md section 0: parity is defined against
    cv2.setUseOptimized(False); a user's OpenCV runs optimized. 80 KITTI-shaped frames (8 seeds x
    10 frames). The survey saw 2 of 80 frames with the same corner set in a different order and
    none with a different set; the bar here is loose on purpose (the number is what matters)."""
    e = engines(1241, 376)
    images = [synth.stereo_frame(seed, t)[0] for seed in range(8) for t in range(0, 40, 4)]
    n, same_set, same_order = default_mode_agreement(e.good_features_to_track, images)
    print(f"\n[default-mode cv2] {n} frames: identical corner set {same_set}/{n}, "
          f"identical order {same_order}/{n}")
    assert same_set >= 0.9 * n and same_order >= 0.8 * n
