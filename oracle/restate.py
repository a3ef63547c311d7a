"""ORACLE (test infrastructure, not product code): numpy restatement of the arithmetic inside
the five OpenCV calls on the reference's hot path, at the reference's literal parameters.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.

The algorithms live in OpenCV, a third-party dependency that is absent from /root/reference
(CMakeLists.txt:27, version unpinned, OpenCV 3.x era). This file restates the published
algorithms of
  cv::goodFeaturesToTrack      (call site src/image_processor.cpp:22)
  cv::StereoBM::compute        (call site src/image_processor.cpp:174-176)
  cv::calcOpticalFlowPyrLK     (call sites src/feature_tracker.cpp:23-26, 32-35)
  the Q-matrix reprojection    (src/image_processor.cpp:178-206)
as plain integer / IEEE-754 arithmetic, with no cv2 call on the math
path. tests/test_oracle_restate.py pins it against cv2 4.13.0 (setUseOptimized(False)) bit for
bit; the CUDA kernels are then written against these formulas. Parity is otherwise unpinned by
the reference (it has no tests or golden vectors).

fl(x) below means "round to float32 after every single operation" (no FMA contraction).
"""
from __future__ import annotations

import math

import numpy as np

F32 = np.float32


def reflect101(i, n):
    """BORDER_REFLECT_101 index: gfedcb|abcdefgh|gfedcba."""
    i = np.asarray(i)
    if n == 1:
        return np.zeros_like(i)
    p = 2 * (n - 1)
    i = np.mod(i, p)
    return np.where(i >= n, p - i, i)


# --------------------------------------------------------------------------------------------
# A1  goodFeaturesToTrack(I, maxCorners, quality, minDistance), blockSize 3, Sobel 3, Shi-Tomasi
# --------------------------------------------------------------------------------------------

def min_eigen_map(img: np.ndarray) -> np.ndarray:
    """cornerMinEigenVal(img, 3, 3) in the SSE3-baseline arithmetic (Appendix A1 steps 1-5)."""
    h, w = img.shape
    scale = 1.0 / (4.0 * 3.0 * 255.0)
    k1 = F32(scale)
    k0 = F32(2.0 * scale)
    ys = reflect101(np.arange(-1, h + 1), h)
    xs = reflect101(np.arange(-1, w + 1), w)
    p = img[np.ix_(ys, xs)].astype(F32)                   # (h+2, w+2), exact small ints
    # Dx: row filter [-1 0 1] (exact), column filter [k1 k0 k1] as fl(fl(d*k0)+fl(fl(up+dn)*k1))
    d = p[:, 2:] - p[:, :-2]                              # exact in f32
    dx = (d[1:-1] * k0).astype(F32) + ((d[:-2] + d[2:]).astype(F32) * k1).astype(F32)
    dx = dx.astype(F32)
    # Dy: row filter [k1 k0 k1] as fl(fl(fl(l*k1)+fl(c*k0))+fl(r*k1)), column filter [-1 0 1]
    s = ((p[:, :-2] * k1).astype(F32) + (p[:, 1:-1] * k0).astype(F32)).astype(F32)
    s = (s + (p[:, 2:] * k1).astype(F32)).astype(F32)
    dy = (s[2:] - s[:-2]).astype(F32)
    xx = (dx * dx).astype(F32)
    xy = (dx * dy).astype(F32)
    yy = (dy * dy).astype(F32)

    def box3(c):
        # row sums in double: (c[x-1] + c[x]) + c[x+1]; columns: sliding accumulator
        cd = c.astype(np.float64)
        cp = cd[:, xs]                                    # (h, w+2)
        r = (cp[:, :-2] + cp[:, 1:-1]) + cp[:, 2:]        # (h, w)
        rp = r[ys, :]                                     # rows -1 .. h
        out = np.empty((h, w), F32)
        acc = (0.0 + rp[0]) + rp[1]
        for y in range(h):
            sacc = acc + rp[y + 2]
            out[y] = sacc.astype(F32)
            acc = sacc - rp[y]
        return out

    a = (box3(xx) * F32(0.5)).astype(F32)
    b = box3(xy)
    c = (box3(yy) * F32(0.5)).astype(F32)
    amc = (a - c).astype(F32)
    t = ((amc * amc).astype(F32) + (b * b).astype(F32)).astype(F32)
    return ((a + c).astype(F32) - np.sqrt(t, dtype=F32)).astype(F32)


def corner_candidates(eig: np.ndarray, quality: float):
    """Steps 6-8: threshold at quality*max, 3x3 dilate-equality NMS on interior pixels, sort by
    value descending with ties broken by larger linear address first. Returns (x, y, val)."""
    h, w = eig.shape
    max_val = eig.max()
    thr = F32(float(max_val) * quality)
    e = np.where(eig > thr, eig, F32(0)).astype(F32)
    pad = np.full((h + 2, w + 2), -np.inf, F32)
    pad[1:-1, 1:-1] = e
    dil = e.copy()
    for dy in range(3):
        for dx in range(3):
            dil = np.maximum(dil, pad[dy:dy + h, dx:dx + w])
    m = (e != 0) & (e == dil)
    m[0, :] = m[-1, :] = False
    m[:, 0] = m[:, -1] = False
    ys, xs = np.nonzero(m)
    vals = e[ys, xs]
    addr = ys.astype(np.int64) * w + xs
    order = np.lexsort((-addr, -vals.astype(np.float64)))
    return xs[order], ys[order], vals[order]


def greedy_min_distance(xs, ys, w, h, max_corners: int, min_distance: float):
    """Steps 9-10: grid-accelerated greedy selection in sorted order."""
    if min_distance < 1:
        n = len(xs) if max_corners <= 0 else min(len(xs), max_corners)
        return np.stack([xs[:n], ys[:n]], 1).astype(F32)
    cell = int(np.rint(min_distance))                     # cvRound
    gw = (w + cell - 1) // cell
    gh = (h + cell - 1) // cell
    grid = [[] for _ in range(gw * gh)]
    md2 = min_distance * min_distance
    out = []
    for x, y in zip(xs.tolist(), ys.tolist()):
        xc, yc = x // cell, y // cell
        x1, y1 = max(0, xc - 1), max(0, yc - 1)
        x2, y2 = min(gw - 1, xc + 1), min(gh - 1, yc + 1)
        good = True
        for yy in range(y1, y2 + 1):
            for xx in range(x1, x2 + 1):
                for (px, py) in grid[yy * gw + xx]:
                    dx, dy = x - px, y - py
                    if dx * dx + dy * dy < md2:
                        good = False
                        break
                if not good:
                    break
            if not good:
                break
        if good:
            grid[yc * gw + xc].append((x, y))
            out.append((x, y))
            if max_corners > 0 and len(out) == max_corners:
                break
    return np.array(out, F32).reshape(-1, 2)


def good_features_to_track(img, max_corners=300, quality=0.1, min_distance=30.0):
    eig = min_eigen_map(img)
    xs, ys, _ = corner_candidates(eig, quality)
    return greedy_min_distance(xs, ys, img.shape[1], img.shape[0], max_corners, min_distance)


# --------------------------------------------------------------------------------------------
# A2  StereoBM(48, 21): XSOBEL prefilter cap 31, texture 10, uniqueness 15, 1/16 px sub-pixel
# --------------------------------------------------------------------------------------------
SBM_NDISP = 48
SBM_WIN = 21
SBM_CAP = 31
SBM_TEXTURE = 10
SBM_UNIQ = 15


def prefilter_xsobel(img: np.ndarray, cap: int = SBM_CAP) -> np.ndarray:
    """Step 1. Rows above/below reflect (101); columns 0 and W-1 are `cap`; when H is odd the
    last row is `cap` (the filter walks rows in pairs)."""
    h, w = img.shape
    im = img.astype(np.int32)
    ys = np.arange(h)
    up = im[reflect101(ys - 1, h)]
    dn = im[reflect101(ys + 1, h)]
    out = np.full((h, w), cap, np.int32)
    d = lambda a: a[:, 2:] - a[:, :-2]
    v = d(up) + 2 * d(im) + d(dn)
    out[:, 1:-1] = np.clip(v, -cap, cap) + cap
    if h % 2 == 1:
        out[h - 1, :] = cap
    return out.astype(np.uint8)


def stereo_bm_valid_roi(w, h, ndisp=SBM_NDISP, win=SBM_WIN):
    """Step 2: x in [ndisp-1 + win/2, w - win/2), y in [win/2, h - win/2)."""
    r = win // 2
    return ndisp - 1 + r, w - r, r, h - r


def stereo_bm_at(pl: np.ndarray, pr: np.ndarray, x: int, y: int, ndisp=SBM_NDISP, win=SBM_WIN,
                 cap=SBM_CAP, texture=SBM_TEXTURE, uniq=SBM_UNIQ) -> int:
    """disp16 of StereoBM::compute at pixel (x, y), evaluated locally (steps 2-7). pl/pr are
    the prefiltered images."""
    h, w = pl.shape
    x0, x1, y0, y1 = stereo_bm_valid_roi(w, h, ndisp, win)
    filtered = -16                                        # (minDisparity - 1) << 4
    if not (x0 <= x < x1 and y0 <= y < y1):
        return filtered
    r = win // 2
    lwin = pl[y - r:y + r + 1, x - r:x + r + 1].astype(np.int32)
    sad = np.empty(ndisp + 2, np.int64)                   # internal index d = ndisp-1-D at [d+1]
    for dd in range(ndisp):
        big_d = ndisp - 1 - dd
        rwin = pr[y - r:y + r + 1, x - r - big_d:x + r + 1 - big_d].astype(np.int32)
        sad[dd + 1] = np.abs(lwin - rwin).sum()
    s = sad[1:ndisp + 1]
    mind = int(np.argmin(s))                              # first minimum
    minsad = int(s[mind])
    tsum = int(np.abs(lwin - cap).sum())
    if tsum < texture:
        return filtered
    if uniq > 0:
        thresh = minsad + (minsad * uniq) // 100
        for dd in range(ndisp):
            if (dd < mind - 1 or dd > mind + 1) and s[dd] <= thresh:
                return filtered
    sad[0] = sad[2]
    sad[ndisp + 1] = sad[ndisp - 1]
    p = int(sad[mind + 2])
    n = int(sad[mind])
    dnm = p + n - 2 * int(sad[mind + 1]) + abs(p - n)
    frac = 0
    if dnm:
        num = (p - n) * 256
        frac = (abs(num) // dnm) * (1 if num >= 0 else -1)   # C division truncates toward zero
    return ((ndisp - mind - 1) * 256 + frac + 15) >> 4


def stereo_bm_sparse(left, right, pts) -> np.ndarray:
    """disp16 at each integer-valued feature (int() truncation as image_processor.cpp:193)."""
    pl = prefilter_xsobel(left)
    pr = prefilter_xsobel(right)
    return np.array([stereo_bm_at(pl, pr, int(px), int(py)) for px, py in pts], np.int16)


# --------------------------------------------------------------------------------------------
# A4  reprojection  w = pose * Q * [x, y, d, 1]^T
# --------------------------------------------------------------------------------------------

def reproject(pts, disp16, pose, focal, cx, cy, baseline):
    """src/image_processor.cpp:178-206. cv::gemm accumulates in double and rounds the result to
    float; returns (valid_idx, pts3d)."""
    focal, cx, cy, baseline = F32(focal), F32(cx), F32(cy), F32(baseline)
    q = np.zeros((4, 4), F32)
    q[0, 0] = q[1, 1] = F32(1.0 / float(focal))
    q[0, 3] = F32(-cx / focal)
    q[1, 3] = F32(-cy / focal)
    q[2, 3] = F32(1)
    q[3, 2] = F32(1.0 / float(F32(baseline * focal)))
    m = (np.asarray(pose, F32).astype(np.float64) @ q.astype(np.float64)).astype(F32)
    idx, out = [], []
    for i, ((x, y), d16) in enumerate(zip(pts, disp16)):
        d = F32(d16) * F32(1.0 / 16)
        if d > 0:
            v = np.array([x, y, d, 1.0], np.float64)
            w = (m.astype(np.float64) @ v).astype(F32)
            idx.append(i)
            out.append((w[0] / w[3], w[1] / w[3], w[2] / w[3]))
    return np.array(idx, np.int32), np.array(out, F32).reshape(-1, 3)


# --------------------------------------------------------------------------------------------
# A3  calcOpticalFlowPyrLK(prev, next, pts, (21,21), 3, (COUNT+EPS, 30, 0.01), 0, 1e-2)
# --------------------------------------------------------------------------------------------

def pyr_down(img: np.ndarray) -> np.ndarray:
    """cv::pyrDown for 8-bit: size ((w+1)/2, (h+1)/2), separable [1 4 6 4 1], reflect-101,
    (sum + 128) >> 8."""
    h, w = img.shape
    oh, ow = (h + 1) // 2, (w + 1) // 2
    im = img.astype(np.int32)
    xs = 2 * np.arange(ow)
    row = sum(k * im[:, reflect101(xs + o, w)] for k, o in zip((1, 4, 6, 4, 1), range(-2, 3)))
    ys = 2 * np.arange(oh)
    col = sum(k * row[reflect101(ys + o, h), :] for k, o in zip((1, 4, 6, 4, 1), range(-2, 3)))
    return ((col + 128) >> 8).astype(np.uint8)


def build_pyramid(img, max_level=3, win=21):
    """Levels as cv::buildOpticalFlowPyramid keeps them (stops when a level is <= win)."""
    levels = [img]
    for _ in range(max_level):
        nxt = pyr_down(levels[-1])
        if nxt.shape[1] <= win or nxt.shape[0] <= win:
            break
        levels.append(nxt)
    return levels


def scharr_deriv(img: np.ndarray):
    """Unnormalised Scharr int16 derivatives with reflect-101 at the image edge."""
    h, w = img.shape
    ys = reflect101(np.arange(-1, h + 1), h)
    xs = reflect101(np.arange(-1, w + 1), w)
    p = img[np.ix_(ys, xs)].astype(np.int32)
    t0 = 3 * (p[:-2] + p[2:]) + 10 * p[1:-1]              # (h, w+2)
    t1 = p[2:] - p[:-2]
    ix = t0[:, 2:] - t0[:, :-2]
    iy = 3 * (t1[:, :-2] + t1[:, 2:]) + 10 * t1[:, 1:-1]
    return ix.astype(np.int16), iy.astype(np.int16)


def _cv_round(v) -> int:
    return int(np.rint(F32(v)))                           # round-half-even, like cvRound (SSE)


def _weights(a, b):
    a, b = F32(a), F32(b)
    one = F32(1.0)
    s = F32(1 << 14)
    iw00 = _cv_round(F32(F32((one - a) * (one - b)) * s))
    iw01 = _cv_round(F32(F32(a * (one - b)) * s))
    iw10 = _cv_round(F32(F32((one - a) * b) * s))
    iw11 = (1 << 14) - iw00 - iw01 - iw10
    return iw00, iw01, iw10, iw11


def _window(arr_pad, pad, ix, iy, n):
    """(n+1)x(n+1) integer window starting at image coordinate (ix, iy) from a padded array."""
    return arr_pad[iy + pad:iy + pad + n + 1, ix + pad:ix + pad + n + 1].astype(np.int64)


def lk_window_sums_a(ixval, iyval):
    """A11, A12, A22 numerators as cv::LKTrackerInvoker accumulates them (lkpyramid.cpp,
    CV_SIMD128 path, no FMA): per window row two 8-pixel SIMD steps, each feeding pixels
    (x0 .. x0+3) then (x0+4 .. x0+7) into four f32 lanes (product rounded, then added), then a
    scalar f32 tail over the remaining pixels 16..20 ((float)(int product), added); at the end
    tail += v_reduce_sum(lanes) = (l0 + l2) + (l1 + l3). Every operation rounds to f32.
    Pinned against cv2 bit for bit by tests/test_oracle_restate.py (through lk_track)."""
    win = ixval.shape[0]
    q11 = np.zeros(4, F32)
    q12 = np.zeros(4, F32)
    q22 = np.zeros(4, F32)
    t11 = t12 = t22 = F32(0)
    simd = (win // 8) * 8
    for y in range(win):
        for x0 in range(0, simd, 4):
            fx = ixval[y, x0:x0 + 4].astype(F32)
            fy = iyval[y, x0:x0 + 4].astype(F32)
            q22 = ((fy * fy).astype(F32) + q22).astype(F32)
            q12 = ((fx * fy).astype(F32) + q12).astype(F32)
            q11 = ((fx * fx).astype(F32) + q11).astype(F32)
        for x in range(simd, win):
            a, b = int(ixval[y, x]), int(iyval[y, x])
            t11 = F32(t11 + F32(a * a))
            t12 = F32(t12 + F32(a * b))
            t22 = F32(t22 + F32(b * b))

    def red(q):
        return F32(F32(q[0] + q[2]) + F32(q[1] + q[3]))
    return F32(t11 + red(q11)), F32(t12 + red(q12)), F32(t22 + red(q22))


def lk_window_sums_b(diff, ixval, iyval):
    """b1, b2 numerators of one LK iteration in cv2's order: per row and 8-pixel SIMD step the
    int32 sums of the pixel pairs (p, p+4), p = 0..3 (v_dotprod), converted to f32 and added to
    lanes qb0 = (p0.x, p0.y, p1.x, p1.y), qb1 = (p2.x, p2.y, p3.x, p3.y); scalar f32 tails over
    pixels 16..20; at the end q = qb0 + qb1, b1 = tail1 + (q0 + q2), b2 = tail2 + (q1 + q3)."""
    win = ixval.shape[0]
    d = diff.astype(np.int64)
    gx = ixval.astype(np.int64)
    gy = iyval.astype(np.int64)
    qb0 = np.zeros(4, F32)
    qb1 = np.zeros(4, F32)
    t1 = t2 = F32(0)
    simd = (win // 8) * 8
    for y in range(win):
        for x0 in range(0, simd, 8):
            px = d[y, x0:x0 + 8] * gx[y, x0:x0 + 8]
            py = d[y, x0:x0 + 8] * gy[y, x0:x0 + 8]
            d0 = np.array([px[0] + px[4], py[0] + py[4], px[1] + px[5], py[1] + py[5]])
            d1 = np.array([px[2] + px[6], py[2] + py[6], px[3] + px[7], py[3] + py[7]])
            qb0 = (qb0 + d0.astype(F32)).astype(F32)
            qb1 = (qb1 + d1.astype(F32)).astype(F32)
        for x in range(simd, win):
            t1 = F32(t1 + F32(int(d[y, x] * gx[y, x])))
            t2 = F32(t2 + F32(int(d[y, x] * gy[y, x])))
    q = (qb0 + qb1).astype(F32)
    return F32(t1 + F32(q[0] + q[2])), F32(t2 + F32(q[1] + q[3]))


def lk_track(prev, nxt, pts, win=21, max_level=3, max_count=30, eps=0.01, min_eig_thr=1e-2,
             exact_sums=False):
    """Per-point pyramidal LK exactly as LKTrackerInvoker walks it (Appendix A3), window sums
    accumulated in cv2's own f32 order (lk_window_sums_a / _b): bit-identical to
    cv2.calcOpticalFlowPyrLK. exact_sums=True sums exact integers instead and converts once
    (what the round-1 kernel did: within 2e-4 px of cv2, but ~20 % of the points differ in the
    last bits). Returns (next_pts float32 (n,2), status uint8 (n,))."""
    pyr_i = build_pyramid(prev, max_level, win)
    pyr_j = build_pyramid(nxt, max_level, win)
    nlev = min(len(pyr_i), len(pyr_j))
    pad = win
    lev = []
    for l in range(nlev):
        im_i, im_j = pyr_i[l], pyr_j[l]
        h, w = im_i.shape
        ys = reflect101(np.arange(-pad, h + pad), h)
        xs = reflect101(np.arange(-pad, w + pad), w)
        ip = im_i[np.ix_(ys, xs)]
        jp = im_j[np.ix_(ys, xs)]
        dx, dy = scharr_deriv(im_i)
        dxp = np.zeros((h + 2 * pad, w + 2 * pad), np.int16)
        dyp = np.zeros_like(dxp)
        dxp[pad:-pad, pad:-pad] = dx
        dyp[pad:-pad, pad:-pad] = dy
        lev.append((ip, jp, dxp, dyp, w, h))
    n = len(pts)
    out = np.zeros((n, 2), F32)
    status = np.ones(n, np.uint8)
    half = F32((win - 1) * 0.5)
    flt_scale = F32(1.0 / (1 << 20))
    eps2 = eps * eps
    for k in range(n):
        npt = np.zeros(2, F32)
        for l in range(nlev - 1, -1, -1):
            ip, jp, dxp, dyp, w, h = lev[l]
            inv = F32(1.0 / (1 << l))
            ppt = np.array([F32(pts[k][0]) * inv, F32(pts[k][1]) * inv], F32)
            if l == nlev - 1:
                npt = ppt.copy()
            else:
                npt = (npt * F32(2.0)).astype(F32)
            out[k] = npt
            ppt = (ppt - half).astype(F32)
            ipx, ipy = int(math.floor(ppt[0])), int(math.floor(ppt[1]))
            if ipx < -win or ipx >= w or ipy < -win or ipy >= h:
                if l == 0:
                    status[k] = 0
                continue
            a = F32(ppt[0] - F32(ipx))
            b = F32(ppt[1] - F32(ipy))
            w00, w01, w10, w11 = _weights(a, b)

            def interp(arr, ix, iy, shift):
                t = _window(arr, pad, ix, iy, win)
                v = t[:-1, :-1] * w00 + t[:-1, 1:] * w01 + t[1:, :-1] * w10 + t[1:, 1:] * w11
                return (v + (1 << (shift - 1))) >> shift

            ival = interp(ip, ipx, ipy, 9)
            ixval = interp(dxp, ipx, ipy, 14)
            iyval = interp(dyp, ipx, ipy, 14)
            if exact_sums:
                s11, s12, s22 = (F32(int((ixval * ixval).sum())), F32(int((ixval * iyval).sum())),
                                 F32(int((iyval * iyval).sum())))
            else:
                s11, s12, s22 = lk_window_sums_a(ixval, iyval)
            a11 = F32(s11 * flt_scale)
            a12 = F32(s12 * flt_scale)
            a22 = F32(s22 * flt_scale)
            d = F32(F32(a11 * a22) - F32(a12 * a12))
            dif = F32(a11 - a22)
            rad = F32(F32(dif * dif) + F32(F32(F32(4.0) * a12) * a12))
            min_eig = F32(F32(F32(a22 + a11) - F32(np.sqrt(rad))) / F32(2 * win * win))
            if float(min_eig) < min_eig_thr or d < np.finfo(F32).eps:
                if l == 0:
                    status[k] = 0
                continue
            d = F32(F32(1.0) / d)
            npt = (npt - half).astype(F32)
            prev_delta = np.zeros(2, F32)
            for j in range(max_count):
                inx, iny = int(math.floor(npt[0])), int(math.floor(npt[1]))
                if inx < -win or inx >= w or iny < -win or iny >= h:
                    if l == 0:
                        status[k] = 0
                    break
                a = F32(npt[0] - F32(inx))
                b = F32(npt[1] - F32(iny))
                w00, w01, w10, w11 = _weights(a, b)
                diff = interp(jp, inx, iny, 9) - ival
                if exact_sums:
                    sb1, sb2 = F32(int((diff * ixval).sum())), F32(int((diff * iyval).sum()))
                else:
                    sb1, sb2 = lk_window_sums_b(diff, ixval, iyval)
                b1 = F32(sb1 * flt_scale)
                b2 = F32(sb2 * flt_scale)
                delta = np.array([F32(F32(F32(a12 * b2) - F32(a22 * b1)) * d),
                                  F32(F32(F32(a12 * b1) - F32(a11 * b2)) * d)], F32)
                npt = (npt + delta).astype(F32)
                out[k] = (npt + half).astype(F32)
                if float(delta[0]) * float(delta[0]) + float(delta[1]) * float(delta[1]) <= eps2:
                    break
                if j > 0 and abs(float(F32(delta[0] + prev_delta[0]))) < 0.01 and \
                        abs(float(F32(delta[1] + prev_delta[1]))) < 0.01:
                    out[k] = (out[k] - (delta * F32(0.5)).astype(F32)).astype(F32)
                    break
                prev_delta = delta
            # the next level restarts from the stored point
            npt = out[k].copy()
        # level-0 epilogue: an `err` vector is passed (src/feature_tracker.cpp:24), so a final
        # bounds check on the window origin can still clear the status
        if status[k]:
            w, h = lev[0][4], lev[0][5]
            fx = F32(out[k][0] - half)
            fy = F32(out[k][1] - half)
            inx, iny = int(math.floor(fx)), int(math.floor(fy))
            if inx < -win or inx >= w or iny < -win or iny >= h:
                status[k] = 0
    return out, status
