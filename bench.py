#!/usr/bin/env python
"""bench.py - stereo frames/s of the front end (detect + pyramid + track fwd/bwd + stereo match +
triangulate) on KITTI-shaped synthetic sequences, one process per GPU.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torchrun)
    python bench.py --impl reference ...                     (the reference's CPU path, host cores:
                                                              its own compiled front end from
                                                              oracle/_ref when that library loads,
                                                              else the cv2 restatement)

Workload (BASELINE configs[3]): 64 independent KITTI-shaped synthetic
stereo sequences per GPU, one texture seed each (rank r owns seeds 64 r .. 64 r + 63), 40
frames each walked forwards and backwards. A "step" is one frame of every sequence of the rank.
The driver loop is stereo_vo_b200.workload.run_steps over svo::BatchFrontEnd - the same loop
tests/test_gpu_bench_path.py checks against the cv2 oracle frame by frame.

Timing: after W warm-up steps the K-step region is repeated R times back to back inside ONE
timed window (CUDA events on every engine stream, max over engine groups and ranks; R chosen so
the window lasts >= --min-seconds, so nvidia-smi clock samples fall inside it). `value` is
measured with the frames already resident in HBM, `e2e` through the C ABI with pinned HOST frames
(H2D of every stereo pair needed and D2H of every result inside the timed window).

Host PnP-RANSAC is outside the accelerated path: a keyframe here keeps all
tracked features and takes the synthetic ground-truth pose; everything the north_star names
(pyramids, corners, stereo matches, temporal tracks, triangulation) runs for every frame.
"""
from __future__ import annotations

import argparse
import csv
import json
import math
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
TOTAL_SEQS_STRONG = 64                         # configs[3]: 64 sequences sharded over the GPUs
# algorithmic bytes per image of the HBM-bound kernels (DESIGN.md 4 / SURVEY.md 8d)
ALGO_BYTES = {
    # fused Shi-Tomasi score + 3x3 NMS: the u8 image is read once, the f32 eig map is never
    # written (SURVEY.md 8d "fused" variant), + 8 B per local maximum (~1e4 per image, ignored)
    "corner_score_nms": N_PIX,
    "pyr_down_u8": 612598 + 153314,            # 3 levels: bytes read + bytes written
}
INGEST_BYTES_PER_IMAGE = 2 * N_PIX             # staged copy read + pitched image written
NCU_SUMMARY = os.path.join(ROOT, "profiles", "r2_ncu_full_summary.csv")
NCU_TRAFFIC = os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")


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

    def __init__(self, gpu_index: int, period_ms: int = 50):
        self.gpu = gpu_index
        self.period_ms = period_ms
        self.proc = None
        self.lines = []
        self.t_lines = []
        self.window = None

    def launch(self):
        """start nvidia-smi early (it needs ~0.2 s to produce its first sample); samples are
        kept only if they fall inside the window set by begin()/end()"""
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.QUERY}",
                 "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())
            self.t_lines.append(time.perf_counter())

    def begin(self):
        self.window = [time.perf_counter(), None]

    def end(self):
        self.window[1] = time.perf_counter()

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"],
                    "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        t0, t1 = self.window if self.window else (0.0, float("inf"))
        for ln, ts in zip(self.lines, self.t_lines):
            if ts < t0 or (t1 is not None and ts > t1):
                continue
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


def measure_h2d_gbs(torch, nbytes=256 << 20, reps=5):
    """pinned host -> device copy rate of this box (one large cudaMemcpyAsync, CUDA events): the
    ceiling of the e2e leg, whose compulsory traffic is one left image per frame + one right image
    per keyframe"""
    src = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    dst = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    best = 0.0
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        dst.copy_(src, non_blocking=True)
        b.record()
        b.synchronize()
        best = max(best, nbytes / (a.elapsed_time(b) * 1e-3) / 1e9)
    del src, dst
    return best


def ncu_matching_figures():
    """SM issue / L1-shared throughput of the matching kernels from the committed ncu capture
    (profiles/, same workload, one engine group) - SURVEY.md 8d asks for these instead of an HBM
    fraction for lk_track / stereo_bm_sparse / sort / greedy."""
    try:
        with open(NCU_SUMMARY) as fh:
            rows = list(csv.DictReader(fh))
    except OSError:
        return None
    out = {}
    for r in rows:
        name = r.get("kernel", "")
        for key in ("lk_track_kernel", "stereo_bm_sparse_48_21_kernel", "sort_small_kernel",
                    "greedy_select_kernel", "stereo_bm_dense_kernel"):
            if name.startswith(key) and key not in out:
                def col(prefix):
                    for k, v in r.items():
                        if k.startswith(prefix):
                            try:
                                return float(v)
                            except ValueError:
                                return None
                    return None
                out[key] = {"issue_active_pct": col("smsp__issue_active"),
                            "l1tex_throughput_pct": col("l1tex__throughput"),
                            "sm_throughput_pct": col("sm__throughput"),
                            "warps_active_pct": col("sm__warps_active"),
                            "registers": col("launch__registers_per_thread")}
    out["source"] = "committed capture " + os.path.relpath(NCU_SUMMARY, ROOT)
    return out


def run_ours(args, rank, world, local_rank):
    import torch
    from stereo_vo_b200 import build, capi, hostapi, sharding, synth, workload
    build.build_lib()                    # no-ops when fresh; serialised across ranks by a file lock
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

    cam = synth.KITTI00
    S, F = args.seqs_per_gpu, args.frames
    seeds = sharding.weak_shard(S, world, rank)          # globally unique texture seeds
    slab = capi.pinned_empty((S, F, 2, H, W), np.uint8)
    t_gen = time.perf_counter()
    workload.generate_frames(seeds, F, slab, workers=max(1, total_cores // max(world, 1)))
    t_gen = time.perf_counter() - t_gen

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x, op="max"):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
        return float(t.item())

    def throughput_leg(n_seq, resident, K, Wm, groups, min_seconds, profile=False,
                       pipeline=True, closed_loop=False):
        """n_seq sequences (the first n_seq of this rank's slab) through hostapi.Batch."""
        batch = hostapi.Batch(W, H, n_seq, groups, cam, device=local_rank, closed_loop=closed_loop)
        if blocking_sync:
            batch.set_blocking_sync(True)
        sub = slab[:n_seq]
        dptr = None
        base = sub.ctypes.data
        if resident:
            dptr = batch.device_alloc(sub.nbytes)
            batch.device_upload(dptr, sub)
            base = dptr
        tables = workload.PointerTables(workload.Schedule(n_seq, F), base, cam)
        pipeline = pipeline and args.pipeline and not profile

        def args_of(k):
            return tables.args(k, W, resident, args.prefetch)

        workload.run_steps(batch, args_of, 0, Wm, pipeline)
        k0 = Wm
        R = 1
        if min_seconds > 0:
            # calibration block (untimed): how many K-step repeats fill the window
            t0 = time.perf_counter()
            workload.run_steps(batch, args_of, k0, k0 + K, pipeline)
            probe = time.perf_counter() - t0
            k0 += K
            R = max(1, int(math.ceil(min_seconds / max(probe, 1e-6))))
            R = int(reduce(float(R)))                     # same repeat count on every rank
        launches0 = batch.kernel_launches
        if profile:
            batch.set_profiling(True)
        sampler = ClockSampler(local_rank)
        sampler.launch()
        if min_seconds > 0:
            time.sleep(0.3)                               # nvidia-smi start-up
        marks = []

        def on_step(k, _sm):
            if (k - k0 + 1) % K == 0:
                marks.append(time.perf_counter())

        barrier()
        sampler.begin()
        batch.timer_begin()
        t0 = time.perf_counter()
        tot = workload.run_steps(batch, args_of, k0, k0 + R * K, pipeline, on_step)
        ev_ms = batch.timer_end()
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        sampler.end()
        clocks = sampler.stop()
        blocks = np.diff(np.array([t0] + marks)) * 1e3 / K      # wall ms per step, per K-block
        out = dict(ms=max(ev_ms, 0.0), wall_ms=wall_ms, clocks=clocks, repeats=R,
                   prof=batch.get_profile() if profile else {},
                   launches=batch.kernel_launches - launches0, tot=tot, n_seq=n_seq,
                   block_ms_per_step=dict(median=float(np.median(blocks)),
                                          min=float(blocks.min()), max=float(blocks.max())))
        if dptr is not None:
            batch.device_free(dptr)
        batch.close()
        return out

    K, Wm = args.steps, args.warmup
    cores_per_rank = max(1, total_cores // max(world, 1))
    if args.groups <= 0:
        # one host worker per engine group + the submitting thread: leave a core for the latter
        args.groups = max(1, min(4, cores_per_rank - 1))
    # spinning waits need a core per thread, which the group count above guarantees; sleeping
    # waits (cudaEventBlockingSync) are only for hosts where even that does not fit. Measured
    # on 8 B200 + 32 cores, 3 groups per rank: spin 445 k frames/s, blocking 253 k (the wake-up
    # latency lands on every step's critical chain)
    blocking_sync = args.sync == "block" or (args.sync == "auto" and cores_per_rank < args.groups + 1)
    dev = throughput_leg(S, True, K, Wm, args.groups, args.min_seconds)
    e2e = dev if args.skip_e2e else throughput_leg(S, False, K, Wm, args.groups, args.min_seconds)
    # closed loop: the pose step (PnP-RANSAC) on the GPU inside the loop instead of teacher poses
    closed = None
    if not args.skip_closed_loop:
        closed = throughput_leg(S, True, K, Wm, args.groups, args.min_seconds / 2, closed_loop=True)
    # strong scaling (configs[3] as written): 64 sequences in total, sharded 64/32/16/8 per GPU
    n_strong = len(sharding.shard_sequences(TOTAL_SEQS_STRONG, world, rank))
    strong = None
    if world > 1 and not args.skip_strong and 0 < n_strong <= S:
        # small per-GPU batches still want every engine group: measured on one B200, 8
        # sequences: 22.3 k frames/s in 4 groups against 19.3 k in 1; 16: 34.0 k against 31.4 k
        strong_groups = min(args.groups, max(1, n_strong // 2))
        strong = throughput_leg(n_strong, True, K, Wm, strong_groups, args.min_seconds / 2)
    # per-kernel durations: a separate pass with ONE engine group and no pipelining, so every
    # kernel is timed alone on the device (with several groups kernels of different streams
    # overlap and their event-bracketed durations include each other)
    Kp = max(1, min(K, args.profile_steps))
    prof_leg = throughput_leg(S, True, Kp, Wm, 1, 0.0, profile=True)

    dev_ms = reduce(dev["ms"])
    e2e_ms = reduce(e2e["ms"])
    dev_frames = world * S * K * dev["repeats"]
    e2e_frames = world * S * K * e2e["repeats"]
    value = dev_frames / (dev_ms * 1e-3)
    e2e_value = e2e_frames / (e2e_ms * 1e-3)
    strong_out = None
    if strong is not None:
        s_ms = reduce(strong["ms"])
        s_frames = reduce(float(strong["n_seq"] * K * strong["repeats"]), "sum")
        strong_out = {"total_sequences": TOTAL_SEQS_STRONG, "seqs_per_gpu": n_strong,
                      "groups_per_gpu": strong_groups,
                      "value": s_frames / (s_ms * 1e-3), "unit": UNIT,
                      "ms_per_step": s_ms / (K * strong["repeats"]), "repeats": strong["repeats"],
                      "note": "frames resident in HBM; same 64 sequences whatever N "
                              "(sharding.shard_sequences), so per-GPU batches shrink with N"}
    closed_out = None
    if closed is not None:
        c_ms = reduce(closed["ms"])
        c_steps = K * closed["repeats"]
        closed_out = {"value": world * S * c_steps / (c_ms * 1e-3), "unit": UNIT,
                      "ms_per_step": c_ms / c_steps, "repeats": closed["repeats"],
                      "keyframes_per_step": closed["tot"]["keyframes"] / c_steps,
                      "note": "frames resident in HBM; keyframe poses from svo_pnp_ransac (GPU "
                              "PnP-RANSAC, cv::solvePnPRansac's arguments of "
                              "src/image_processor.cpp:80) on the tracked features and the stored "
                              "world points: svo_step -> svo_pnp_ransac -> svo_keyframe -> "
                              "svo_tracker_set, no caller poses"}
    h2d_gbs = measure_h2d_gbs(torch) if rank == 0 else None

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
                images = S * Kp + prof_leg["tot"]["keyframes"]
                gbs = INGEST_BYTES_PER_IMAGE * images / (ms * 1e-3) / 1e9
                ent.update(algo_gbs=gbs, frac_of_hbm=gbs / peak)
            kernels[name] = ent
        # the dominant kernel of a step by time: the two LK launches are one kernel
        by_kernel = {}
        for name, ent in kernels.items():
            base = "lk_track" if name.startswith("lk_track") else name
            by_kernel[base] = by_kernel.get(base, 0.0) + ent["ms_per_step"]
        dom = max(by_kernel.items(), key=lambda kv: kv[1])[0] if by_kernel else None
        # roofline of the dominant kernel among those whose interface is HBM traffic (whole-image
        # passes); lk_track / stereo / sort / greedy work on <= 400 points per frame and are
        # issue/latency-bound: their SM issue and L1 figures are under `matching`
        hbm_dom = max((n for n in ("corner_score_nms", "pyr_down_u8") if n in kernels),
                      key=lambda n: kernels[n]["ms_per_step"], default=None)
        traffic, traffic_src = None, None
        try:
            with open(NCU_TRAFFIC) as fh:
                ncu_t = json.load(fh)
            pick = [k for k in ncu_t if k.startswith(
                {"corner_score_nms": "corner_score_nms_kernel",
                 "pyr_down_u8": "pyr_down"}.get(hbm_dom, "\0"))]
            if pick:
                traffic = sum(ncu_t[k]["dram_bytes_per_launch"] for k in pick)
                traffic_src = ("committed capture " + os.path.relpath(NCU_TRAFFIC, ROOT) +
                               " (dram__bytes_read.sum + dram__bytes_write.sum per launch; not "
                               "re-measured in this run)")
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
                                "segment (DESIGN.md 4); HBM-bound kernels: ingest, pyr_down_u8 "
                                "(see `kernels`)",
                        "peak_source": peak_src,
                        "algorithmic_bytes_per_launch": ALGO_BYTES[hbm_dom] * S,
                        "dominant_kernel_of_step": dom,
                        "dominant_kernel_ms_per_step": by_kernel.get(dom),
                        "matching": ncu_matching_figures()}
        cpu = None if args.skip_cpu_baseline else cpu_baseline(sample_frames=args.cpu_frames)
        # every left image + the right images actually transferred (predicted keyframes and
        # keyframes fetched on demand; counted by the driver)
        e2e_steps = K * e2e["repeats"]
        rights_per_step = e2e["tot"]["rights"] / e2e_steps
        in_bytes = int(S * H * W + rights_per_step * H * W)
        compulsory = int(S * H * W + e2e["tot"]["keyframes"] / e2e_steps * H * W)
        eng_tmp_block = 16 + 24 + 300 * 8 + 400 * 16
        e2e_step_ms = e2e_ms / e2e_steps
        pcie = None
        if h2d_gbs:
            pcie = {"h2d_gbs_measured": h2d_gbs,
                    "h2d_gbs_achieved": in_bytes / (e2e_step_ms * 1e-3) / 1e9,
                    "frac_of_link": in_bytes / (e2e_step_ms * 1e-3) / 1e9 / h2d_gbs,
                    "compulsory_bytes_per_step": compulsory,
                    "ceiling_frames_per_s": h2d_gbs * 1e9 / (compulsory / S),
                    "frac_of_ceiling": (S / (e2e_step_ms * 1e-3)) / (h2d_gbs * 1e9 / (compulsory / S)),
                    "note": "ceiling = measured H2D rate / compulsory bytes per frame (every frame "
                            "must deliver its left image, every keyframe its right image); "
                            + ("the e2e leg is link-bound" if in_bytes / (e2e_step_ms * 1e-3) / 1e9 / h2d_gbs > 0.75
                               else "the e2e leg is NOT link-bound: copies run under the kernels, the "
                                    "gap to `value` is the ingest of host frames + on-demand right images")}
        extras = None
        if world == 1 and not args.skip_extras:
            extras = run_extras(local_rank)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": Wm, "ms_per_step": dev_ms / (K * dev["repeats"]), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "u8/f32/f64 (bit-exact OpenCV arithmetic)", "data": "synthetic",
            "repeats": dev["repeats"], "timed_steps": K * dev["repeats"],
            "timed_region_s": dev_ms * 1e-3,
            "block_ms_per_step_wall": dev["block_ms_per_step"],
            "config": {"workload": f"{S} independent KITTI-shaped (1241x376) synthetic stereo "
                                   f"sequences per GPU (texture seeds {S}*rank .. {S}*rank+{S - 1}, "
                                   f"{F} frames each, ping-pong), kitti00 params "
                                   "(BASELINE configs[3] shape; configs[1]/[2]/[4] under `extras`)",
                       "seqs_per_gpu": S, "frames_per_step": S * world,
                       "frames_per_sequence": F,
                       "parallelism": f"{world} GPU(s) x {S} sequences in {args.groups} engine groups "
                                      "(one CUDA stream + host worker each), no collective",
                       "host_cores": total_cores, "rank0_cpu_binding": numa_bind,
                       "gpu_wait": "blocking (cudaEventBlockingSync)" if blocking_sync else "spin",
                       "host": "C++ svo::BatchFrontEnd (libsvo_host.so) via "
                               + ("svo_batch_submit/wait, 2 steps in flight" if args.pipeline
                                  else "svo_batch_step")
                               + "; right images uploaded for keyframes only; e2e leg "
                               + ("prefetches" if args.prefetch else "does not prefetch")
                               + " the next step's frames (svo_prefetch)",
                       "l2": f"inputs larger than L2: {S * F * 2 * N_PIX / 1e9:.2f} GB of frames "
                             f"per GPU, every step reads {S} new stereo pairs "
                             f"({S * 2 * N_PIX / 1e6:.0f} MB) plus pyramids; no flush between steps",
                       "timing": f"K = {K} steps repeated R = {dev['repeats']} times back to back in "
                                 "one CUDA-event window (max over engine groups and ranks)",
                       "keyframes_per_step": dev["tot"]["keyframes"] / (K * dev["repeats"]),
                       "tracked_per_frame": dev["tot"]["tracked"] / (S * K * dev["repeats"]),
                       "corners_per_frame": dev["tot"]["corners"] / (S * K * dev["repeats"]),
                       "frame_generation_s": t_gen,
                       "pnp": "host PnP outside the accelerated path: keyframes keep all tracks, "
                              "pose = synthetic ground truth",
                       "parity": "this exact driver loop is compared with the cv2 oracle in "
                                 "tests/test_gpu_bench_path.py (64 sequences x 56 steps, both legs)",
                       "reference_arm": "--impl reference runs the cv2 restatement of the "
                                        "reference, one sequence per host core x --ref-frames "
                                        "frames per step (a bounded sample: the CPU needs ~25 ms "
                                        "per frame); same metric and frame shape, so the two arms' "
                                        "frames/s compare directly although steps/config differ"},
            "clocks": dev["clocks"],
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_step_ms,
                    "repeats": e2e["repeats"], "timed_region_s": e2e_ms * 1e-3,
                    "h2d_bytes_per_step": in_bytes,
                    "d2h_bytes_per_step": S * eng_tmp_block,
                    "block_ms_per_step_wall": e2e["block_ms_per_step"],
                    "pcie": pcie,
                    "clocks": e2e["clocks"]},
            "gpu_launches": dev["launches"],
            "gpu_launches_per_step": dev["launches"] / (K * dev["repeats"]),
            "roofline": roofline,
            "kernels": kernels,
            "closed_loop": closed_out,
            "strong_scaling": strong_out,
            "extras": extras,
            "cpu_baseline": cpu,
            "wall_ms_per_step": dev["wall_ms"] / (K * dev["repeats"]),
        }
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return line


def run_extras(device: int) -> dict:
    """The other BASELINE configs, measured in the same run (N = 1 only): configs[1]/[2]
    single-sequence latency (KITTI, d435i), configs[4] 3840x2160 / 20 k corners throughput, and
    the dense StereoBM map. All through the same C++ driver / C ABI."""
    from stereo_vo_b200 import capi, hostapi, synth, workload
    out = {}
    # ---- single-sequence latency: one synchronous svo_batch_step per frame, host frames
    for name, cam, motion in (("kitti_latency", synth.KITTI00, synth.KITTI_MOTION),
                              ("d435i_latency", synth.D435I, synth.D435I_MOTION)):
        n_fr = 40
        slab = capi.pinned_empty((1, n_fr, 2, cam.height, cam.width), np.uint8)
        workload.generate_frames([3], n_fr, slab, cam, motion, workers=1)
        batch = hostapi.Batch(cam.width, cam.height, 1, 1, cam, device=device)
        tables = workload.PointerTables(workload.Schedule(1, n_fr), slab.ctypes.data, cam, motion)
        lat, kinds = [], []
        for k in range(3 * n_fr):
            a = tables.args(k, cam.width, False, False)
            t0 = time.perf_counter()
            sm = batch.step_raw(*a)
            dt = (time.perf_counter() - t0) * 1e3
            if k >= 8:
                lat.append(dt)
                kinds.append(sm.keyframes > 0)
        batch.close()
        lat, kf = np.array(lat), np.array(kinds)
        out[name] = {"config": f"{cam.width}x{cam.height}, one sequence, synchronous svo_batch_step "
                               "per frame (H2D + kernels + D2H + host bookkeeping), wall clock",
                     "ms_per_frame_median": float(np.median(lat)),
                     "ms_p95": float(np.percentile(lat, 95)),
                     "ms_skip_frames": float(np.median(lat[~kf])) if (~kf).any() else None,
                     "ms_keyframes": float(np.median(lat[kf])) if kf.any() else None,
                     "frames": int(len(lat)), "keyframes": int(kf.sum())}
    # ---- configs[4]: 3840x2160, goodFeaturesToTrack(20000, 0.01, 20), feature limit lifted
    try:
        cam, motion = synth.UHD4K, synth.UHD4K_MOTION
        S4, F4 = 4, 4
        slab = capi.pinned_empty((S4, F4, 2, cam.height, cam.width), np.uint8)
        workload.generate_frames(list(range(S4)), F4, slab, cam, motion)
        p = capi.default_params(cam.width, cam.height)
        p.max_corners, p.quality_level, p.min_distance, p.max_features = 20000, 0.01, 20.0, 20000
        batch = hostapi.Batch(cam.width, cam.height, S4, 2, cam, device=device, params=p)
        dptr = batch.device_alloc(slab.nbytes)
        batch.device_upload(dptr, slab)
        tables = workload.PointerTables(workload.Schedule(S4, F4, phase_step=1), dptr, cam, motion)
        workload.run_steps(batch, lambda k: tables.args(k, cam.width, True, False), 0, 4, True)
        n_steps = 24
        batch.timer_begin()
        tot = workload.run_steps(batch, lambda k: tables.args(k, cam.width, True, False), 4,
                                 4 + n_steps, True)
        ms = batch.timer_end()
        batch.device_free(dptr)
        batch.close()
        out["uhd_stress"] = {"config": f"{S4} sequences of 3840x2160 resident in HBM "
                                       f"({slab.nbytes / 1e6:.0f} MB of frames > L2), "
                                       "goodFeaturesToTrack(20000, 0.01, 20), max_features 20000, "
                                       "4-level pyramid, 2 engine groups, pipelined",
                             "frames_per_s": S4 * n_steps / (ms * 1e-3),
                             "ms_per_frame": ms / (S4 * n_steps),
                             "corners_per_frame": tot["corners"] / (S4 * n_steps),
                             "tracked_per_frame": tot["tracked"] / (S4 * n_steps),
                             "keyframes": tot["keyframes"]}
    except Exception as ex:                                # report, never hide the headline
        out["uhd_stress"] = {"error": repr(ex)}
    # ---- dense StereoBM maps per second, 64 KITTI pairs resident on the device
    try:
        cam = synth.KITTI00
        n = 64
        e = capi.Engine(cam.width, cam.height, n_seq=n, device=device)
        pairs = [synth.stereo_frame(s, 0) for s in range(8)]
        for s in range(n):
            e.stereo_bm_dense(pairs[s % 8][0], pairs[s % 8][1], seq=s)
        seqs = list(range(n))
        for _ in range(3):
            e.stereo_bm_dense_resident(seqs, fetch=False)
        ms = []
        for _ in range(20):
            e.stereo_bm_dense_resident(seqs, fetch=False)
            ms.append(e.last_timing()[1])
        e.close()
        ms = float(np.median(ms))
        out["dense_stereo_bm"] = {"config": f"{n} KITTI pairs resident on the device, one launch, "
                                            "full CV_16S map of StereoBM(48, 21)",
                                  "maps_per_s": n / (ms * 1e-3), "us_per_pair": 1e3 * ms / n,
                                  "algo_gbs": 4 * N_PIX * n / (ms * 1e-3) / 1e9}
    except Exception as ex:
        out["dense_stereo_bm"] = {"error": repr(ex)}
    return out


CPU_STAGES = ("goodFeaturesToTrack", "calcOpticalFlowPyrLK", "StereoBM.compute", "solvePnPRansac")


def _time_cv2_stages(cv2, acc):
    """Wrap the four OpenCV entry points the reference calls (oracle/cv2_frontend.py looks them
    up on the cv2 module at call time) so that their wall time lands in acc[name]."""
    def timed(name, fn):
        def wrapper(*a, **k):
            t = time.perf_counter()
            try:
                return fn(*a, **k)
            finally:
                acc[name] += time.perf_counter() - t
        return wrapper

    for name in ("goodFeaturesToTrack", "calcOpticalFlowPyrLK", "solvePnPRansac"):
        setattr(cv2, name, timed(name, getattr(cv2, name)))
    create = cv2.StereoBM_create

    class TimedStereoBM:
        def __init__(self, *a, **k):
            self._bm = create(*a, **k)
            self.compute = timed("StereoBM.compute", self._bm.compute)

        def __getattr__(self, item):
            return getattr(self._bm, item)

    cv2.StereoBM_create = TimedStereoBM


def _cpu_worker(arg):
    seed, n_frames, threads, optimized, impl = arg
    import cv2
    from oracle import cv2_frontend as O
    from stereo_vo_b200 import synth
    cv2.setNumThreads(threads)
    # True: the reference's realistic speed (default OpenCV mode); False: the canonical
    # arithmetic the parity tests compare with
    cv2.setUseOptimized(bool(optimized))
    stages = {k: 0.0 for k in CPU_STAGES}
    _time_cv2_stages(cv2, stages)
    # sequences of at most 80 frames each: on longer synthetic runs tracked + new features pass
    # the 400-feature limit (from frame ~89 on), where the reference's own bookkeeping is
    # undefined behaviour and the oracle refuses to continue
    chunks, left = [], n_frames
    while left > 0:
        c = min(80, left)
        chunks.append(synth.sequence(seed + 1000 * len(chunks), c))
        left -= c
    if impl == "compiled":
        # the reference's own compiled process() / track_features() / triangulate_stereo() /
        # add_keyframe() (oracle/_ref/libref_frontend.so), its OpenCV calls served by this
        # process's cv2 through callbacks
        from oracle import ref_frontend
        f, cx, cy, b = synth.KITTI00.f32()
        t0 = time.perf_counter()
        done = 0
        for fr in chunks:
            ref = ref_frontend.RefPipeline(f, cx, cy, b, guess="as_stored", record=False)
            for t, (l, r) in enumerate(fr):
                ref.process(l, r, t / 11.0)
                done += 1
            ref.close()
        return time.perf_counter() - t0, max(done, 1), stages
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
    return time.perf_counter() - t0, max(done, 1), stages


def cpu_run(n_proc: int, frames_each: int, threads_each: int, optimized: bool = True, impl: str = "port"):
    """The reference's CPU front end on the host cores: n_proc independent sequences in parallel
    processes. impl "port": the cv2 restatement (oracle/cv2_frontend.py); "compiled": the
    reference's own sources compiled where they lie (oracle/_ref). Returns aggregate frames/s,
    the slowest process's wall time, the caller's wall time and the mean ms per frame spent
    inside each of the reference's OpenCV calls (+ "other": the rest of process())."""
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    args = [(100 + i, frames_each, threads_each, optimized, impl) for i in range(n_proc)]
    t0 = time.perf_counter()
    if n_proc == 1 and impl == "port":
        res = [_cpu_worker(args[0])]
        wall = res[0][0]
    else:
        with ctx.Pool(n_proc) as pool:
            # (bounded: a worker that died would leave map() waiting for ever)
            res = pool.map_async(_cpu_worker, args).get(timeout=900)
        wall = max(r[0] for r in res)
    total = sum(r[1] for r in res)
    stage_ms = {k: 1e3 * sum(r[2][k] for r in res) / total for k in CPU_STAGES}
    stage_ms["other"] = 1e3 * sum(r[0] for r in res) / total - sum(stage_ms.values())
    return total / wall, wall, time.perf_counter() - t0, stage_ms


def reference_impl():
    """("compiled", None) when oracle/_ref/libref_frontend.so is usable on this machine, else
    ("port", why). Checked by running two frames through it in a child process."""
    if os.environ.get("SVO_BENCH_REF_IMPL") == "port":
        return "port", "SVO_BENCH_REF_IMPL=port"
    try:
        fps, _, _, _ = cpu_run(1, 2, 1, True, impl="compiled")
        return ("compiled", None) if fps > 0 else ("port", "no frames")
    except Exception as ex:                      # noqa: BLE001
        return "port", repr(ex)[:160]


_KIND = {"compiled": "reference", "port": "port"}
_WHAT = {"compiled": "oracle/_ref/libref_frontend.so = the reference's own image_processor.cpp + "
                     "feature_tracker.cpp + bundle_adjuster.cpp compiled where they lie (OpenCV's "
                     "algorithms served by cv2 through callbacks, Ceres replaced by a recording "
                     "stand-in: no solve)",
         "port": "oracle/cv2_frontend.py = the reference restated call for call on cv2"}


def cpu_baseline(sample_frames: int = 120):
    cores = len(os.sched_getaffinity(0))
    n_proc = max(1, min(cores, 32))
    each = max(8, sample_frames // n_proc) if n_proc > 1 else sample_frames
    impl, why_port = reference_impl()
    try:
        fps, wall, _, stage_ms = cpu_run(n_proc, each, 1, impl=impl)
    except Exception as ex:                      # noqa: BLE001
        if impl == "port":
            raise
        impl, why_port = "port", repr(ex)[:160]
        fps, wall, _, stage_ms = cpu_run(n_proc, each, 1, impl=impl)
    # the canonical (setUseOptimized(False)) arithmetic the parity tests compare with, on a
    # quarter of the sample: SURVEY.md 8d asks for both modes
    c_each = max(8, each // 4)
    c_fps, c_wall, _, c_stage_ms = cpu_run(n_proc, c_each, 1, optimized=False, impl=impl)
    rnd = lambda d: {k: round(v, 3) for k, v in d.items()}
    out = {"value": fps, "unit": UNIT, "cores": n_proc, "kind": _KIND[impl],
           "sample": f"{n_proc} processes x {each} synthetic KITTI-shaped frames, {_WHAT[impl]}; cv2 "
                     f"{_cv2_version()}, 1 OpenCV thread each, default optimized mode; "
                     f"{wall:.1f} s wall = {wall * n_proc:.0f} core-seconds of CPU work",
           "stage_ms_per_frame": rnd(stage_ms),
           "stage_note": "mean wall ms per frame and core inside each OpenCV call of the "
                         "reference (calcOpticalFlowPyrLK = forward + backward; StereoBM.compute "
                         "and solvePnPRansac run on keyframes only, amortised over all frames)",
           "canonical_mode": {"value": c_fps, "unit": UNIT,
                              "sample": f"{n_proc} processes x {c_each} frames, "
                                        f"cv2.setUseOptimized(False), {c_wall:.1f} s wall",
                              "stage_ms_per_frame": rnd(c_stage_ms)}}
    if impl == "compiled":
        # the Python restatement beside it (the oracle of the parity tests), a quarter of the sample
        p_fps, p_wall, _, _ = cpu_run(n_proc, c_each, 1, impl="port")
        out["port"] = {"value": p_fps, "unit": UNIT, "kind": "port",
                       "sample": f"{n_proc} processes x {c_each} frames, {_WHAT['port']}, {p_wall:.1f} s wall"}
    else:
        out["compiled_reference_unavailable"] = why_port
    return out


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
    impl, why_port = reference_impl()

    def timed_steps(which):
        for _ in range(args.warmup if args.warmup < 2 else 1):
            cpu_run(n_proc, max(4, each // 2), 1, impl=which)
        t_sum, n_sum = 0.0, 0
        for _ in range(K):
            _fps, wall, _, _stages = cpu_run(n_proc, each, 1, impl=which)
            t_sum += wall
            n_sum += n_proc * each
        return t_sum, n_sum

    K = max(1, min(args.steps, 3))
    try:
        t_total, frames_total = timed_steps(impl)
    except Exception as ex:                      # noqa: BLE001
        if impl == "port":
            raise
        impl, why_port = "port", repr(ex)[:160]
        t_total, frames_total = timed_steps(impl)
    value = frames_total / t_total
    sample = (f"{n_proc} processes x {each} synthetic KITTI-shaped frames per step, {_WHAT[impl]}; cv2 "
              f"{_cv2_version()}" + ("" if impl == "compiled" else f" (compiled reference not usable here: {why_port})"))
    return {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
        "steps": K, "warmup": args.warmup, "ms_per_step": t_total / K * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/f32/f64",
        "data": "synthetic",
        "config": {"workload": f"{n_proc} independent KITTI-shaped (1241x376) synthetic stereo "
                               "sequences on the host cores, kitti00 params",
                   "note": "bounded sample of the GPU arm's workload: one sequence per host core, "
                           f"{each} frames per step, at most 3 steps (the CPU path needs ~25 ms "
                           "per frame and core); frames/s is aggregate over all cores and "
                           "compares directly with the GPU arm's frames/s"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": n_proc, "kind": _KIND[impl],
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--seqs-per-gpu", type=int, default=64)
    ap.add_argument("--frames", type=int, default=40, help="frames per synthetic sequence")
    ap.add_argument("--min-seconds", type=float, default=2.0,
                    help="minimum length of the timed window of each leg: the K-step region is "
                         "repeated back to back until it lasts this long")
    ap.add_argument("--cpu-frames", type=int, default=2400)
    ap.add_argument("--skip-cpu-baseline", action="store_true", help="profiling runs only")
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs only")
    ap.add_argument("--skip-strong", action="store_true", help="no strong-scaling leg (N > 1)")
    ap.add_argument("--skip-closed-loop", action="store_true", help="no closed-loop (GPU PnP) leg")
    ap.add_argument("--skip-extras", action="store_true",
                    help="no d435i / 4K / dense extras (they run at N = 1 only)")
    ap.add_argument("--ref-frames", type=int, default=16)
    ap.add_argument("--profile-steps", type=int, default=60,
                    help="steps of the separate single-group pass that times every kernel alone")
    ap.add_argument("--groups", type=int, default=0,
                    help="engine groups (CUDA streams + host workers) the sequences of a GPU are "
                         "split into; 0 = 4, or fewer when the ranks of this node have less than "
                         "4 host cores each")
    ap.add_argument("--sync", default="auto", choices=["auto", "spin", "block"],
                    help="how engine calls wait for the GPU: spin (lowest latency), block (sleep on an "
                         "interrupt), auto = block when a rank has fewer host cores than threads")
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
