"""The multi-sequence workload of BASELINE configs[3]: S independent
KITTI-shaped synthetic stereo sequences per GPU, one texture seed each, and the loop that drives
them through svo::BatchFrontEnd. bench.py times this loop; tests/test_gpu_bench_path.py runs the
very same loop and compares every per-frame output with the cv2 oracle, so the path the
headline number is measured on is the path that is parity-checked.

Host-side plumbing only (frame generation, pointer tables, the submit/wait order); no pixel
work happens here.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Callable, List, Optional, Sequence

import numpy as np

from . import synth


def _gen_sequence(arg):
    seed, n_frames, cam, motion = arg
    out = np.empty((n_frames, 2, cam.height, cam.width), np.uint8)
    for t in range(n_frames):
        l, r = synth.stereo_frame(seed, t, cam, motion)
        out[t, 0] = l
        out[t, 1] = r
    return seed, out


def generate_frames(seeds: Sequence[int], n_frames: int, out: np.ndarray,
                    cam: synth.Camera = synth.KITTI00, motion: synth.Motion = synth.KITTI_MOTION,
                    workers: Optional[int] = None) -> None:
    """Fill out[slot, t, 0/1] (uint8 [S][F][2][H][W], e.g. pinned memory) with frames 0..F-1 of
    sequence seeds[slot]. One worker process per sequence at a time (spawned: the parent may
    already hold a CUDA context)."""
    jobs = [(int(s), n_frames, cam, motion) for s in seeds]
    if workers is None:
        workers = len(os.sched_getaffinity(0))
    workers = max(1, min(workers, len(jobs)))
    slot_of = {}
    for i, s in enumerate(seeds):
        slot_of.setdefault(int(s), []).append(i)
    if workers == 1:
        results = map(_gen_sequence, jobs)
        for seed, arr in results:
            for i in slot_of[seed]:
                out[i] = arr
        return
    import multiprocessing as mp
    with mp.get_context("spawn").Pool(workers) as pool:
        for seed, arr in pool.imap_unordered(_gen_sequence, jobs):
            for i in slot_of[seed]:
                out[i] = arr


class Schedule:
    """Which frame of its sequence a slot sees at step k. Every slot walks its F frames
    forwards, then backwards (ping-pong: consecutive steps are always consecutive frames, so the
    tracker sees real inter-frame motion however long the run is), starting `phase_step` * slot
    steps into that cycle: the sequences of a GPU are independent camera streams, so their
    keyframes do not fall on the same steps."""

    def __init__(self, n_slots: int, n_frames: int, phase_step: int = 5):
        self.n_slots = n_slots
        self.n_frames = n_frames
        self.period = max(1, 2 * n_frames - 2)
        self.phase = [(phase_step * s) % self.period for s in range(n_slots)]

    def frame(self, step: int, slot: int) -> int:
        k = (step + self.phase[slot]) % self.period
        return k if k < self.n_frames else 2 * self.n_frames - 2 - k


def pose_of(t: int, motion: synth.Motion = synth.KITTI_MOTION) -> np.ndarray:
    """camera -> world pose handed to a keyframe at frame t: synthetic ground truth, R = I."""
    p = np.eye(4, dtype=np.float32)
    p[:3, 3] = synth.ground_truth(t, motion)
    return p


class PointerTables:
    """ctypes pointer arrays (left, right, poses) for every step of the periodic schedule over a
    frame slab at address `base` laid out [S][F][2][H][W] (host or device memory)."""

    def __init__(self, sched: Schedule, base: int, cam: synth.Camera = synth.KITTI00,
                 motion: synth.Motion = synth.KITTI_MOTION):
        self.sched = sched
        S, F = sched.n_slots, sched.n_frames
        img = cam.height * cam.width
        self.rows = []
        for k in range(sched.period):
            ls, rs, ps = [], [], []
            for slot in range(S):
                t = sched.frame(k, slot)
                off = ((slot * F + t) * 2) * img
                ls.append(base + off)
                rs.append(base + off + img)
                ps.append(pose_of(t, motion))
            poses = np.ascontiguousarray(np.stack(ps).reshape(S, 16), np.float32)
            self.rows.append(((C.c_void_p * S)(*ls), (C.c_void_p * S)(*rs), poses,
                              poses.ctypes.data_as(C.c_void_p)))

    def args(self, k: int, pitch: int, device: bool, prefetch: bool):
        """positional arguments of hostapi.Batch.step_raw / submit_raw for step k"""
        lp, rp, _poses, pz = self.rows[k % self.sched.period]
        if device or not prefetch:
            return (lp, rp, pitch, device, pz)
        nlp, nrp, _p, _z = self.rows[(k + 1) % self.sched.period]
        return (lp, rp, pitch, False, pz, nlp, nrp)


def run_steps(batch, args_of: Callable[[int], tuple], first: int, last: int, pipeline: bool,
              on_step: Optional[Callable[[int, object], None]] = None) -> dict:
    """Steps first..last-1 through `batch` (hostapi.Batch). pipeline: two steps in flight
    (svo_batch_submit / svo_batch_wait), else one synchronous svo_batch_step per step.
    on_step(k, summary) runs after step k completed (its results are the ones batch.get reads).
    Returns the summed counters."""
    tot = dict(tracked=0, corners=0, keyframes=0, rights=0, frames=0)

    def fold(k, sm):
        tot["tracked"] += sm.tracked
        tot["corners"] += sm.corners
        tot["keyframes"] += sm.keyframes
        tot["rights"] += sm.right_images
        tot["frames"] += sm.frames
        if on_step is not None:
            on_step(k, sm)

    if last <= first:
        return tot
    if pipeline:
        batch.submit_raw(*args_of(first))
        for k in range(first, last):
            if k + 1 < last:
                batch.submit_raw(*args_of(k + 1))
            fold(k, batch.wait())
    else:
        for k in range(first, last):
            fold(k, batch.step_raw(*args_of(k)))
    return tot
