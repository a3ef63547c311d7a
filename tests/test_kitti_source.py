"""CPU: the KITTI frame source (src/kitti_node.cpp:16-68 without ROS): pose file and 8-bit PNG
pairs read by libsvo_host.so must equal what the reference's own calls give - cv2.imread(...,
IMREAD_GRAYSCALE) for the images (:67-68) and the 12 numbers per line for the poses (:47-50).
The PNGs are written with cv2 (zlib-compressed, adaptive filters) into a KITTI-shaped tree."""
import os

import cv2
import numpy as np
import pytest

from stereo_vo_b200 import synth
from tests import util


@pytest.fixture(scope="module")
def kitti_tree(tmp_path_factory):
    root = tmp_path_factory.mktemp("kitti")
    seq = "07"
    cam = util.small_camera(421, 127)                       # odd sizes: row filters see every phase
    os.makedirs(root / seq / "image_0")
    os.makedirs(root / seq / "image_1")
    os.makedirs(root / "data_odometry_poses" / "dataset" / "poses")
    n = 7
    poses = []
    rng = np.random.default_rng(4)
    frames = []
    for i in range(n):
        l, r = synth.stereo_frame(2, i, cam)
        if i == 3:                                          # an RGB file: imread converts to grey
            cv2.imwrite(str(root / seq / "image_0" / f"{i:06d}.png"), np.dstack([l, r, l]))
        else:
            cv2.imwrite(str(root / seq / "image_0" / f"{i:06d}.png"), l,
                        [cv2.IMWRITE_PNG_COMPRESSION, 1 + i])
        cv2.imwrite(str(root / seq / "image_1" / f"{i:06d}.png"), r)
        frames.append((l, r))
        p = np.eye(4)[:3].copy()
        p[:, 3] = rng.normal(size=3) * i
        p[:3, :3] += rng.normal(scale=1e-3, size=(3, 3))
        poses.append(p)
    with open(root / "data_odometry_poses" / "dataset" / "poses" / f"{seq}.txt", "w") as fh:
        for p in poses:
            fh.write(" ".join(f"{v:.12e}" for v in p.reshape(-1)) + "\n")
        fh.write("1 2 3\n")                                 # a short line ends the sequence (:47)
    return str(root) + "/", seq, n, poses, cam


def test_frames_and_poses_match_the_reference_reader(kitti_tree):
    from stereo_vo_b200 import build, hostapi
    if build.host_needs_build():
        build.build_host_lib()
    root, seq, n, poses, cam = kitti_tree
    src = hostapi.KittiSource(root, seq)
    assert (src.n_frames, src.width, src.height) == (n, cam.width, cam.height)
    for i in range(n):
        l, r, p = src.read(i)
        ref_l = cv2.imread(f"{root}{seq}/image_0/{i:06d}.png", cv2.IMREAD_GRAYSCALE)
        ref_r = cv2.imread(f"{root}{seq}/image_1/{i:06d}.png", cv2.IMREAD_GRAYSCALE)
        assert np.array_equal(l, ref_l), f"left {i}"
        assert np.array_equal(r, ref_r), f"right {i}"
        assert np.allclose(p, poses[i], rtol=0, atol=1e-11)
    from stereo_vo_b200.capi import SvoError
    with pytest.raises(SvoError, match="past the pose file"):
        src.read(n)
    src.close()
    with pytest.raises(SvoError, match="cannot open"):
        hostapi.KittiSource(root, "99")


def _write_png(path, img, filters):
    """A PNG writer for the tests: 8-bit grey / grey+alpha / RGB / RGBA, row y filtered with
    type filters[y % len(filters)] (0 None, 1 Sub, 2 Up, 3 Average, 4 Paeth), IDAT split in two."""
    import struct
    import zlib
    h, w = img.shape[:2]
    ch = 1 if img.ndim == 2 else img.shape[2]
    color = {1: 0, 2: 4, 3: 2, 4: 6}[ch]
    rows = img.reshape(h, w * ch).astype(np.int32)
    prev = np.zeros(w * ch, np.int32)
    raw = bytearray()
    for y in range(h):
        f = filters[y % len(filters)]
        cur = rows[y]
        a = np.concatenate([np.zeros(ch, np.int32), cur[:-ch]])
        c = np.concatenate([np.zeros(ch, np.int32), prev[:-ch]])
        b = prev
        if f == 0:
            pred = np.zeros_like(cur)
        elif f == 1:
            pred = a
        elif f == 2:
            pred = b
        elif f == 3:
            pred = (a + b) >> 1
        else:
            p = a + b - c
            pa, pb, pc = np.abs(p - a), np.abs(p - b), np.abs(p - c)
            pred = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, b, c))
        raw.append(f)
        raw += ((cur - pred) & 255).astype(np.uint8).tobytes()
        prev = cur
    z = zlib.compress(bytes(raw), 6)

    def chunk(kind, data):
        return struct.pack(">I", len(data)) + kind + data + struct.pack(">I", zlib.crc32(kind + data))

    with open(path, "wb") as fh:
        fh.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, color, 0, 0, 0))
                 + chunk(b"tEXt", b"Comment\0test") + chunk(b"IDAT", z[:len(z) // 2])
                 + chunk(b"IDAT", z[len(z) // 2:]) + chunk(b"IEND", b""))


@pytest.mark.parametrize("channels", [1, 2, 3, 4])
def test_every_png_filter_type_and_colour_type(kitti_tree, tmp_path, channels):
    """cv2's PNG writer only emits Sub-filtered rows; KITTI's files use all five filter types.
    Frames written with every filter type in turn (and one file per single type), as grey,
    grey + alpha, RGB and RGBA, must decode to exactly what cv2.imread(..., GRAYSCALE) returns."""
    import shutil
    from stereo_vo_b200 import build, hostapi
    if build.host_needs_build():
        build.build_host_lib()
    root, seq, n, poses, cam = kitti_tree
    base = tmp_path / "tree"
    shutil.copytree(root, base)
    base = str(base) + "/"
    rng = np.random.default_rng(channels)
    plans = [[0, 1, 2, 3, 4], [4, 3, 2, 1, 0, 4, 4, 3], [0], [1], [2], [3], [4]]
    for i, filters in enumerate(plans):
        l, r = synth.stereo_frame(5, i, cam)
        if channels > 1:
            extra = [rng.integers(0, 256, l.shape, dtype=np.uint8) for _ in range(channels - 1)]
            l = np.dstack([l] + extra)
            r = np.dstack([r] + extra[::-1])
        _write_png(f"{base}{seq}/image_0/{i:06d}.png", l, filters)
        _write_png(f"{base}{seq}/image_1/{i:06d}.png", r, filters[::-1])
    src = hostapi.KittiSource(base, seq)
    for i in range(len(plans)):
        l, r, _ = src.read(i)
        ref_l = cv2.imread(f"{base}{seq}/image_0/{i:06d}.png", cv2.IMREAD_GRAYSCALE)
        ref_r = cv2.imread(f"{base}{seq}/image_1/{i:06d}.png", cv2.IMREAD_GRAYSCALE)
        assert ref_l is not None and ref_l.shape == (cam.height, cam.width)
        assert np.array_equal(l, ref_l), f"left {i} (filters {plans[i]})"
        assert np.array_equal(r, ref_r), f"right {i}"
    src.close()


def test_malformed_pngs_are_refused_with_a_message(kitti_tree, tmp_path):
    """A frame file that is not a sound 8-bit PNG of the sequence's size must end in an error
    that names the file - never in a crash or a silently wrong image."""
    import shutil
    import struct
    import zlib
    from stereo_vo_b200 import build, hostapi
    from stereo_vo_b200.capi import SvoError
    if build.host_needs_build():
        build.build_host_lib()
    root, seq, n, poses, cam = kitti_tree
    good = open(f"{root}{seq}/image_0/000000.png", "rb").read()

    def tree_with(first_left: bytes):
        base = tmp_path / f"t{len(os.listdir(tmp_path))}"
        shutil.copytree(root, base)
        with open(base / seq / "image_0" / "000000.png", "wb") as fh:
            fh.write(first_left)
        return str(base) + "/"

    def chunk(kind: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + kind + data + struct.pack(">I", zlib.crc32(kind + data))

    def png(w, h, depth=8, color=0, payload=None):
        raw = payload if payload is not None else zlib.compress(b"\0" * ((w + 1) * h))
        ihdr = struct.pack(">IIBBBBB", w & 0xffffffff, h & 0xffffffff, depth, color, 0, 0, 0)
        return b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", ihdr) + chunk(b"IDAT", raw) + chunk(b"IEND", b"")

    # the size comes from the first left frame: implausible headers stop the constructor
    for bad_file, what in [(good[:20], "not a PNG"), (b"JFIF" + good[4:], "not a PNG"),
                           (png(0, 5), "implausible"), (png(-3, 5), "implausible"),
                           (png(1 << 20, 5), "implausible")]:
        with pytest.raises(SvoError, match=what):
            hostapi.KittiSource(tree_with(bad_file), seq)
    # a sound header with an unsound body fails at the read
    for bad_file, what in [(good[:60], "truncated"), (png(cam.width, cam.height, depth=16), "8-bit"),
                           (png(cam.width, cam.height, color=3), "palette"),
                           (png(cam.width, cam.height, payload=b"garbage"), "inflate"),
                           (png(cam.width, cam.height,
                                payload=zlib.compress(b"\7" * ((cam.width + 1) * cam.height))), "filter"),
                           (png(cam.width, cam.height - 1), "size differs")]:
        base = tree_with(good)
        src = hostapi.KittiSource(base, seq)
        with open(f"{base}{seq}/image_1/000002.png", "wb") as fh:
            fh.write(bad_file)
        with pytest.raises(SvoError, match=what):
            src.read(2)
        src.close()


@pytest.mark.parametrize("threads,depth", [(1, 3), (3, 5), (6, 8), (4, 3)])
def test_prefetcher_threads_deliver_every_frame_once_in_order(kitti_tree, threads, depth):
    """Several decoder threads, frames finishing out of order: the consumer still sees 0, 1, 2, ...
    with the bytes of a direct read, each in the slot index % depth, and a slot keeps its frame
    until the second further next() (two steps may be in flight: frames i and i - 1)."""
    import time
    from stereo_vo_b200 import build, hostapi
    if build.host_needs_build():
        build.build_host_lib()
    root, seq, n, poses, cam = kitti_tree
    direct = hostapi.KittiSource(root, seq)
    want = [direct.read(i) for i in range(n)]
    for first, slow in [(0, False), (2, True)]:
        src = hostapi.KittiSource(root, seq)
        seen, held = [], []
        for i, l, r, p in src.prefetch(depth=depth, first=first, threads=threads, pinned=False):
            if slow:
                time.sleep(0.02)                   # the ring fills up; workers must wait
            assert np.array_equal(l, want[i][0]) and np.array_equal(r, want[i][1]), f"frame {i}"
            assert np.allclose(p, want[i][2], rtol=0, atol=1e-11)
            held.append((i, l, r))
            for j, hl, hr in held[-2:]:            # this frame and the one before it are intact
                assert np.array_equal(hl, want[j][0]) and np.array_equal(hr, want[j][1]), (i, j)
            seen.append(i)
        assert seen == list(range(first, n))
        src.close()
    direct.close()


def test_prefetcher_reports_a_bad_frame_after_the_good_ones(kitti_tree, tmp_path):
    import shutil
    from stereo_vo_b200 import build, hostapi
    from stereo_vo_b200.capi import SvoError
    if build.host_needs_build():
        build.build_host_lib()
    root, seq, n, poses, cam = kitti_tree
    base = tmp_path / "tree"
    shutil.copytree(root, base)
    base = str(base) + "/"
    with open(f"{base}{seq}/image_1/000004.png", "r+b") as fh:      # corrupt the zlib stream
        fh.seek(200)
        fh.write(b"\xff" * 64)
    src = hostapi.KittiSource(base, seq)
    seen = []
    with pytest.raises(SvoError, match="000004.png"):
        for i, l, r, p in src.prefetch(depth=6, threads=3, pinned=False):
            seen.append(i)
    assert seen == [0, 1, 2, 3]
    src.close()
    # closing while the workers are still decoding ahead must not hang or crash
    for _ in range(3):
        src = hostapi.KittiSource(root, seq)
        it = src.prefetch(depth=5, threads=4, pinned=False)
        next(it)
        src.close()


@pytest.mark.gpu
def test_decode_ahead_ring_feeds_the_front_end(kitti_tree, svo_lib):
    """The prefetcher thread decodes into a pinned ring while the GPU works on earlier frames:
    every frame arrives once, in order, with the bytes of a direct read, and goes straight into
    svo_step from the ring slot."""
    from stereo_vo_b200 import hostapi
    from stereo_vo_b200.capi import Engine
    root, seq, n, poses, cam = kitti_tree
    src = hostapi.KittiSource(root, seq)
    direct = hostapi.KittiSource(root, seq)
    e = Engine(cam.width, cam.height, n_seq=1)
    seen = []
    prev = None
    for i, l, r, p in src.prefetch(depth=3):
        dl, dr, dp = direct.read(i)
        assert np.array_equal(l, dl) and np.array_equal(r, dr) and np.allclose(p, dp)
        if prev is not None:                     # the previous slot is still intact
            assert np.array_equal(prev[1], direct.read(prev[0])[0])
        res = e.step([0], [l], [r])
        assert res["n_corners"][0] > 4
        seen.append(i)
        prev = (i, l)
    assert seen == list(range(n))
    e.close()
    src.close()
    direct.close()
