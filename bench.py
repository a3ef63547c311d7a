#!/usr/bin/env python
"""bench.py - stereo frames/s of the front end (detect + pyramid + track fwd/bwd + stereo match +
triangulate) on KITTI-shaped synthetic sequences, one process per GPU.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torchrun)
    python bench.py --impl reference ...                     (the reference's CPU path, host cores)

A "step" is one frame index of every sequence resident on the GPU (S stereo frames per rank;
workload: SURVEY.md 8d config (4), 64 independent KITTI-shaped sequences per GPU, weak scaling:
sequences are independent, no collective on the data path). The JSON line follows the driver's
contract; `value` is measured with frames already resident in HBM, `e2e` through the C ABI with
pinned HOST buffers (H2D of every stereo pair and D2H of every result inside the timed region).

Host PnP-RANSAC is outside the accelerated path: a keyframe here keeps all
tracked features and takes the synthetic ground-truth pose; everything the north_star names
(pyramids, corners, stereo matches, temporal tracks, triangulation) runs for every frame.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "stereo frames/s (front end, 1241x376)"
UNIT = "frames/s"
W, H = 1241, 376
N_PIX = W * H
# algorithmic bytes per image of the HBM-bound kernels (DESIGN.md / SURVEY.md 8d)
ALGO_BYTES = {
    # fused Shi-Tomasi score + 3x3 NMS: the u8 image is read once, the f32 eig map is never
    # written (SURVEY.md 8d "fused" variant), + 8 B per local maximum (~1e4 per image, ignored)
    "corner_score_nms": N_PIX,
    "pyr_down_u8": 612598 + 153314,            # 3 levels: bytes read + bytes written
}
INGEST_BYTES_PER_IMAGE = 2 * N_PIX             # staged copy read + pitched image written


def _env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.QUERY}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[5 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(max(smax)) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_frames(n_tex: int, n_frames: int):
    """n_tex distinct synthetic sequences x n_frames stereo pairs (pinned host memory)."""
    from stereo_vo_b200 import capi, synth
    buf = capi.pinned_empty((n_tex, n_frames, 2, H, W), np.uint8)
    for s in range(n_tex):
        for t in range(n_frames):
            l, r = synth.stereo_frame(s, t)
            buf[s, t, 0] = l
            buf[s, t, 1] = r
    return buf


def frame_index(step: int, slot: int, n_tex: int, n_frames: int):
    """Sequence slot -> (texture, frame). Slots sharing a texture run at different phases; the
    frame index ping-pongs so consecutive frames of a slot are always consecutive in time."""
    tex = slot % n_tex
    phase = (slot // n_tex) * 3
    k = (step + phase) % (2 * n_frames - 2)
    t = k if k < n_frames else 2 * n_frames - 2 - k
    return tex, t


def bind_near_gpu(local_rank: int):
    """Keep this rank (and the pinned host frames it first touches) on the host cores / NUMA node
    its GPU hangs off: with 8 ranks on a two-socket box the H2D stream of every rank otherwise
    crosses the socket link half of the time. Best effort: returns what it did, or None."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = torch.cuda.get_device_properties(local_rank).pci_domain_id
        dev = torch.cuda.get_device_properties(local_rank).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/local_cpulist"
        with open(path) as fh:
            txt = fh.read().strip()
        cpus = set()
        for part in txt.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if not cpus or cpus == os.sched_getaffinity(0):
            return None
        os.sched_setaffinity(0, cpus)
        return f"cpus {txt}"
    except Exception:
        return None


def run_ours(args, rank, world, local_rank):
    import torch
    from stereo_vo_b200 import capi, frontend, synth
    from stereo_vo_b200 import build
    if build.needs_build():
        build.build_lib()
    if build.host_needs_build():
        build.build_host_lib()
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the front end has no CPU path")
    torch.cuda.set_device(local_rank)
    total_cores = len(os.sched_getaffinity(0))
    numa_bind = bind_near_gpu(local_rank) if world > 1 and not args.no_bind else None
    dist = None
    if world > 1:
        # NCCL_DEBUG=VERSION makes NCCL print its version banner on stdout, next to the one JSON
        # line the driver parses
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    S = args.seqs_per_gpu
    n_tex, n_frames = args.textures, args.frames
    frames = make_frames(n_tex, n_frames)
    cam = synth.KITTI00

    def pose_fn(slot, step):
        tex, t = frame_index(step, slot, n_tex, n_frames)
        p = np.eye(4, dtype=np.float32)
        p[:3, 3] = synth.ground_truth(t)
        return p

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def run_leg(device_resident: bool, steps: int, warmup: int, profile: bool, groups: int):
        import ctypes as C
        from stereo_vo_b200 import hostapi
        batch = hostapi.Batch(W, H, S, groups, cam, device=local_rank)
        pitch = W
        dptr = None
        if device_resident:
            dptr = batch.device_alloc(frames.nbytes)
            batch.device_upload(dptr, frames)
        img_bytes = H * W
        base = dptr if device_resident else frames.ctypes.data
        period = 2 * n_frames - 2
        # pointer arrays and keyframe poses for every phase of the (periodic) frame schedule
        sched = []
        for k in range(period):
            ls, rs, ps = [], [], []
            for slot in range(S):
                tex, t = frame_index(k, slot, n_tex, n_frames)
                off = ((tex * n_frames + t) * 2) * img_bytes
                ls.append(base + off)
                rs.append(base + off + img_bytes)
                ps.append(pose_fn(slot, k))
            poses = np.ascontiguousarray(np.stack(ps).reshape(S, 16), np.float32)
            sched.append(((C.c_void_p * S)(*ls), (C.c_void_p * S)(*rs), poses,
                          poses.ctypes.data_as(C.c_void_p)))

        def args_of(k):
            lp, rp, _poses, pz = sched[k % period]
            if device_resident or not args.prefetch:
                return (lp, rp, pitch, device_resident, pz)
            nlp, nrp, _p, _z = sched[(k + 1) % period]
            return (lp, rp, pitch, False, pz, nlp, nrp)

        def one(k):
            return batch.step_raw(*args_of(k))

        for k in range(warmup):
            one(k)
        launches0 = batch.kernel_launches
        if profile:
            batch.set_profiling(True)
        sampler = ClockSampler(local_rank)
        barrier()
        sampler.start()
        batch.timer_begin()
        t0 = time.perf_counter()
        tracked = corners = keyframes = rights = 0
        first, last = warmup, warmup + steps
        if args.pipeline and not profile:
            # two steps in flight (svo_batch_submit / svo_batch_wait): engine groups run free of
            # each other, the tail of step k overlaps the heavy kernels of step k+1
            batch.submit_raw(*args_of(first))
            for k in range(first, last):
                if k + 1 < last:
                    batch.submit_raw(*args_of(k + 1))
                sm = batch.wait()
                tracked += sm.tracked
                corners += sm.corners
                keyframes += sm.keyframes
                rights += sm.right_images
        else:
            for k in range(first, last):
                sm = one(k)
                tracked += sm.tracked
                corners += sm.corners
                keyframes += sm.keyframes
                rights += sm.right_images
        ev_ms = batch.timer_end()
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        clocks = sampler.stop()
        prof = batch.get_profile() if profile else {}
        out = dict(ms=max(ev_ms, 0.0), wall_ms=wall_ms, clocks=clocks, prof=prof,
                   launches=batch.kernel_launches - launches0, keyframes=keyframes,
                   tracked=tracked, corners=corners, rights=rights)
        if dptr is not None:
            batch.device_free(dptr)
        batch.close()
        return out

    K, Wm = args.steps, args.warmup
    if args.groups <= 0:
        cores_per_rank = max(1, total_cores // max(world, 1))
        args.groups = min(4, cores_per_rank)
    dev = run_leg(True, K, Wm, profile=False, groups=args.groups)
    e2e = dev if args.skip_e2e else run_leg(False, K, Wm, profile=False, groups=args.groups)
    # per-kernel durations: a separate pass with ONE engine group, so every kernel is timed
    # alone on the device (with several groups kernels of different streams overlap and their
    # event-bracketed durations include each other)
    Kp = max(1, min(K, args.profile_steps))
    prof_leg = run_leg(True, Kp, Wm, profile=True, groups=1)

    def reduce_max(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    dev_ms = reduce_max(dev["ms"])
    e2e_ms = reduce_max(e2e["ms"])
    total_frames = world * S * K
    value = total_frames / (dev_ms * 1e-3)
    e2e_value = total_frames / (e2e_ms * 1e-3)

    line = None
    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
                peaks = json.load(fh)
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else \
            "fallback 6650 GB/s (B200_PROFILING.md)"
        prof = prof_leg["prof"]
        kernels = {}
        step_kernel_ms = sum(ms for ms, _ in prof.values())
        for name, (ms, cnt) in prof.items():
            if cnt == 0:
                continue
            ent = {"ms_per_step": ms / Kp, "share": ms / step_kernel_ms if step_kernel_ms else None}
            if name in ALGO_BYTES:
                gbs = ALGO_BYTES[name] * S * cnt / (ms * 1e-3) / 1e9
                ent.update(algo_gbs=gbs, frac_of_hbm=gbs / peak)
            elif name == "ingest":
                # every left image + the right image of every keyframe (svo_set_right)
                images = S * Kp + prof_leg["keyframes"]
                gbs = INGEST_BYTES_PER_IMAGE * images / (ms * 1e-3) / 1e9
                ent.update(algo_gbs=gbs, frac_of_hbm=gbs / peak)
            kernels[name] = ent
        dom = max(prof.items(), key=lambda kv: kv[1][0])[0] if prof else None
        # roofline of the dominant kernel among those whose interface is HBM traffic (whole-image
        # passes); lk_track / stereo / sort / greedy work on <= 400 points per frame and are
        # issue/latency-bound (profiles/: SM issue and L1 figures from ncu)
        hbm_dom = max((n for n in ("corner_score_nms", "pyr_down_u8") if n in kernels),
                      key=lambda n: kernels[n]["ms_per_step"], default=None)
        # DRAM traffic per launch of that kernel from the committed `ncu --set full` capture
        # (profiles/, same command line, one engine group); None when no capture is on disk
        traffic, traffic_src = None, None
        try:
            with open(os.path.join(ROOT, "profiles", "r1_ncu_traffic.json")) as fh:
                ncu_t = json.load(fh)
            cs_name = next((k for k in ncu_t if k.startswith("corner_score_nms_kernel")),
                           "corner_score_nms_kernel")
            pick = {"corner_score_nms": [cs_name],
                    "pyr_down_u8": ["pyr_down_u8_kernel<8>", "pyr_down_u8_kernel<4>",
                                    "pyr_down_u8_kernel<2>"]}.get(hbm_dom, [])
            if pick and all(k in ncu_t for k in pick):
                traffic = sum(ncu_t[k]["dram_bytes_per_launch"] for k in pick)
                traffic_src = "profiles/r1_ncu_traffic.json (dram__bytes_read.sum + dram__bytes_write.sum)"
        except (OSError, ValueError, KeyError):
            pass
        roofline = None
        if hbm_dom:
            roofline = {"bound": "hbm", "kernel": hbm_dom,
                        "achieved": kernels[hbm_dom]["algo_gbs"], "peak": peak, "unit": "GB/s",
                        "frac": kernels[hbm_dom]["frac_of_hbm"], "traffic": traffic,
                        "traffic_source": traffic_src,
                        "note": "corner_score_nms is issue-bound, not HBM-bound: OpenCV's sliding "
                                "fp64 column sums force one thread per image column and row "
                                "segment (DESIGN.md 4; segments whose guessed start was wrong are "
                                "re-walked by `corner_repair`, timed separately in `kernels`); "
                                "HBM-bound kernels: ingest, pyr_down_u8 (see `kernels`)",
                        "peak_source": peak_src,
                        "algorithmic_bytes_per_launch": ALGO_BYTES[hbm_dom] * S,
                        "dominant_kernel_of_step": dom}
        cpu = None if args.skip_cpu_baseline else cpu_baseline(sample_frames=args.cpu_frames)
        # every left image + the right images actually transferred (predicted keyframes and
        # keyframes fetched on demand; counted by the driver)
        in_bytes = int(S * H * W + e2e["rights"] / K * H * W)
        eng_tmp_block = 16 + 24 + 300 * 8 + 400 * 16
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": Wm, "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8/f32/f64 (bit-exact OpenCV arithmetic)",
            "data": "synthetic",
            "config": {"workload": f"{S} independent KITTI-shaped (1241x376) synthetic stereo "
                                   "sequences per GPU, kitti00 params (BASELINE configs[3] shape)",
                       "seqs_per_gpu": S, "frames_per_step": S * world,
                       "parallelism": f"{world} GPU(s) x {S} sequences in {args.groups} engine groups "
                                      "(one CUDA stream + host worker each), no collective",
                       "host_cores": total_cores, "rank0_cpu_binding": numa_bind,
                       "host": "C++ svo::BatchFrontEnd (libsvo_host.so) via "
                               + ("svo_batch_submit/wait, 2 steps in flight" if args.pipeline
                                  else "svo_batch_step")
                               + "; right images uploaded for keyframes only; e2e leg "
                               + ("prefetches" if args.prefetch else "does not prefetch")
                               + " the next step's frames (svo_prefetch)",
                       "l2": "inputs larger than L2: every step reads 64 new stereo pairs (60 MB "
                             "staged + 60 MB pitched copies + 2 x 51 MB pyramid slots in flight "
                             "> 126 MB L2); no flush between steps",
                       "keyframes_per_step": dev["keyframes"] / K,
                       "tracked_per_frame": dev["tracked"] / (S * K),
                       "corners_per_frame": dev["corners"] / (S * K),
                       "pnp": "host PnP outside the accelerated path: keyframes keep all tracks, "
                              "pose = synthetic ground truth"},
            "clocks": dev["clocks"],
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_ms / K,
                    "h2d_bytes_per_step": in_bytes,
                    "d2h_bytes_per_step": S * eng_tmp_block,
                    "clocks": e2e["clocks"]},
            "gpu_launches": dev["launches"],
            "roofline": roofline,
            "kernels": kernels,
            "cpu_baseline": cpu,
            "wall_ms_per_step": dev["wall_ms"] / K,
        }
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return line


def _cpu_worker(arg):
    seed, n_frames, threads = arg
    import cv2
    from oracle import cv2_frontend as O
    from stereo_vo_b200 import synth
    cv2.setNumThreads(threads)
    cv2.setUseOptimized(True)             # the reference's realistic speed (default OpenCV mode)
    # sequences of at most 48 frames each: on longer synthetic runs tracked + new features pass
    # the 400-feature limit, where the reference's own bookkeeping is undefined behaviour
    #  and the oracle refuses to continue
    chunks, left = [], n_frames
    while left > 0:
        c = min(48, left)
        chunks.append(synth.sequence(seed + 1000 * len(chunks), c))
        left -= c
    t0 = time.perf_counter()
    done = 0
    for fr in chunks:
        proc = O.make_pipeline(synth.KITTI00)
        try:
            for l, r in fr:
                proc.process(l, r)
                done += 1
        except AssertionError:            # B6 reached early on this seed: next sequence
            continue
    return time.perf_counter() - t0, max(done, 1)


def cpu_run(n_proc: int, frames_each: int, threads_each: int):
    """The reference's CPU front end (cv2 restatement, oracle/) on the host cores: n_proc
    independent sequences in parallel processes. Returns aggregate frames/s."""
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    args = [(100 + i, frames_each, threads_each) for i in range(n_proc)]
    t0 = time.perf_counter()
    if n_proc == 1:
        res = [_cpu_worker(args[0])]
        wall = res[0][0]
    else:
        with ctx.Pool(n_proc) as pool:
            res = pool.map(_cpu_worker, args)
        wall = max(r[0] for r in res)
    total = sum(r[1] for r in res)
    return total / wall, wall, time.perf_counter() - t0


def cpu_baseline(sample_frames: int = 120):
    cores = len(os.sched_getaffinity(0))
    n_proc = max(1, min(cores, 32))
    each = max(8, sample_frames // n_proc) if n_proc > 1 else sample_frames
    fps, wall, _ = cpu_run(n_proc, each, 1)
    return {"value": fps, "unit": UNIT, "cores": n_proc, "kind": "port",
            "sample": f"{n_proc} processes x {each} synthetic KITTI-shaped frames, cv2 "
                      f"{_cv2_version()} oracle (oracle/cv2_frontend.py: the reference's own "
                      "OpenCV calls), 1 OpenCV thread each, default optimized mode; "
                      f"{wall:.1f} s wall = {wall * n_proc:.0f} core-seconds of CPU work"}


def _cv2_version():
    try:
        import cv2
        return cv2.__version__
    except Exception:
        return "?"


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on the host cores."""
    if rank != 0:
        return None
    cores = len(os.sched_getaffinity(0))
    n_proc = max(1, min(cores, 64))
    each = args.ref_frames
    vals = []
    for _ in range(args.warmup if args.warmup < 2 else 1):
        cpu_run(n_proc, max(4, each // 2), 1)
    t_total = 0.0
    frames_total = 0
    K = max(1, min(args.steps, 3))
    for _ in range(K):
        fps, wall, _ = cpu_run(n_proc, each, 1)
        vals.append(fps)
        t_total += wall
        frames_total += n_proc * each
    value = frames_total / t_total
    sample = (f"{n_proc} processes x {each} synthetic KITTI-shaped frames per step, cv2 "
              f"{_cv2_version()} oracle restatement of src/image_processor.cpp + "
              "src/feature_tracker.cpp (reference not buildable here: ROS/OpenCV SDK/Ceres absent)")
    return {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
        "steps": K, "warmup": args.warmup, "ms_per_step": t_total / K * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/f32/f64",
        "data": "synthetic",
        "config": {"workload": f"{n_proc} independent KITTI-shaped (1241x376) synthetic stereo "
                               "sequences on the host cores, kitti00 params"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": n_proc, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--seqs-per-gpu", type=int, default=64)
    ap.add_argument("--textures", type=int, default=8)
    ap.add_argument("--frames", type=int, default=12)
    ap.add_argument("--cpu-frames", type=int, default=2400)
    ap.add_argument("--skip-cpu-baseline", action="store_true", help="profiling runs only")
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs only")
    ap.add_argument("--ref-frames", type=int, default=16)
    ap.add_argument("--profile-steps", type=int, default=60,
                    help="steps of the separate single-group pass that times every kernel alone")
    ap.add_argument("--groups", type=int, default=0,
                    help="engine groups (CUDA streams + host workers) the sequences of a GPU are "
                         "split into; 0 = 4, or fewer when the ranks of this node have less than "
                         "4 host cores each (every worker waits for its stream on a core)")
    ap.add_argument("--no-bind", action="store_true",
                    help="multi-GPU runs: do not pin the rank to the host cores next to its GPU")
    ap.add_argument("--no-pipeline", dest="pipeline", action="store_false",
                    help="one synchronous svo_batch_step per step instead of two steps in flight")
    ap.add_argument("--no-prefetch", dest="prefetch", action="store_false",
                    help="e2e leg: do not copy the next step's frames ahead")
    args = ap.parse_args()
    rank = _env_int("RANK", 0)
    world = _env_int("WORLD_SIZE", 1)
    local_rank = _env_int("LOCAL_RANK", 0)
    if world == 1 and args.gpus > 1:
        # launched without torchrun: run N independent replicas is not possible in one process
        print(f"[bench] --gpus {args.gpus} without torchrun: running rank 0 only", file=sys.stderr)
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        line = run_reference(args, rank, world)
    else:
        line = run_ours(args, rank, world, local_rank)
    if rank == 0 and line is not None:
        print(json.dumps(line))


if __name__ == "__main__":
    main()
