#!/usr/bin/env python
"""bench.py -- frames/s of the iterative 3D Hough hot path on BASELINE.json config[1]:
single ToF frames 224x172 (38 528 px) -> PassThrough + VoxelGrid (leaf 0.05/1.5) -> Hough at
granularity 6, opt_nlines 20, opt_minvotes 12, dx = sqrt(3)*leaf.

  python bench.py --gpus N --steps K --warmup W            our arm (libh3d.so, CUDA)
  python bench.py --impl reference --gpus N --steps K ...  CPU arm: the reference's own code
                                                           (oracle/_ref, else the oracle port)

A step is one pass of the hot path over a batch of FRAMES_PER_STEP distinct synthetic frames
(poses along the tower trajectory).  `value` is timed with the frames resident in HBM
(h3d_pipeline_run_dev), `e2e` through h3d_pipeline_run with pinned HOST frames: H2D of the raw
frame and D2H of the line records and inlier lists are inside the timed region.  Times are CUDA
event times on the pipeline's stream (the library records them around every frame), summed over
the frames of the timed steps; for N > 1 every rank replays its own frames (frame-parallel,
no collective on the data path) and the maximum over ranks is reported.
Only the cpu_baseline leg and --impl reference execute anything under oracle/.
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
sys.path.insert(0, ROOT)

FRAMES_PER_STEP = 16
GRANULARITY = 6
NLINES = 20
MINVOTES = 12
METRIC = "frames/s (38k-pt ToF frame: PassThrough+VoxelGrid, Hough gran 6, opt_nlines 20)"
UNIT = "frames/s"
L2_FLUSH_BYTES = 256 << 20


def make_frames(count, rank=0):
    """`count` distinct raw ToF frames (38 528 x 4 f32) along the tower trajectory, cached in /tmp."""
    from pointcloud_segmentation_b200 import synth
    cache = f"/tmp/h3d_frames_{count}_{rank}.npy"
    if os.path.exists(cache):
        f = np.load(cache)
        if f.shape == (count, synth.TOF_W * synth.TOF_H, 4):
            return f
    poses = synth.tower_trajectory(1000)
    pick = np.linspace(0, 999, count).astype(int) if rank == 0 else (np.arange(count) * 37 + 101 * rank) % 1000
    frames = np.stack([synth.tof_frame(pose=tuple(poses[k]), seed=1000 + int(k)) for k in pick]).astype(np.float32)
    try:
        np.save(cache, frames)
    except OSError:
        pass
    return frames


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [s for s in sm if s > 0.5 * max(mx or [1])] or sm
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------
# CPU arm
# ---------------------------------------------------------------------------------------------
def _cpu_worker(args):
    """One frame through the CPU path: pre-filter restatement + the reference library's loop."""
    frame, use_ref = args
    from oracle import oracle as O
    from pointcloud_segmentation_b200 import synth
    orc = O.Oracle()
    eng = O.Reference() if use_ref else orc
    t0 = time.perf_counter()
    filt = orc.prefilter(frame, synth.WINDOW, float(synth.LEAF))
    tr = eng.hough3dlines(filt.astype(np.float64), GRANULARITY, synth.OPT_DX, MINVOTES, NLINES, mode=0)
    return time.perf_counter() - t0, tr.n_lines


def _cpu_init(use_ref):
    # the reference builds its direction sphere in O(V^2) (15 s at g6): do it once per worker,
    # outside the timed region, like the node does at start-up (node.cpp:929)
    from oracle import oracle as O
    if use_ref:
        O.Reference().lib.ref_init_space(GRANULARITY)


def cpu_reference_available():
    from oracle import oracle as O
    return O.Reference.available()


def run_cpu_arm(args):
    """--impl reference: the reference's CPU implementation, one process per host core (the
    reference is single-threaded and frames are independent), each step = one frame per worker."""
    import multiprocessing as mp
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as O
    O.build(ref=None)
    use_ref = cpu_reference_available()
    cores = os.cpu_count() or 1
    workers = max(1, min(cores, 32))
    frames = make_frames(FRAMES_PER_STEP)
    with mp.get_context("spawn").Pool(workers, initializer=_cpu_init, initargs=(use_ref,)) as pool:
        jobs = [(frames[k % len(frames)], use_ref) for k in range(workers)]
        for _ in range(args.warmup):
            pool.map(_cpu_worker, jobs)
        t0 = time.perf_counter()
        lines = 0
        for _ in range(args.steps):
            lines += sum(r[1] for r in pool.map(_cpu_worker, jobs))
        dt = time.perf_counter() - t0
    value = workers * args.steps / dt
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "configs[1]: single 224x172 ToF frames, g6, opt_nlines 20", "frames_per_step": workers,
                   "lines_per_frame": lines / max(1, workers * args.steps)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": "reference" if use_ref else "port",
                         "sample": f"{workers} frames per step, one single-threaded process per core"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(out), flush=True)


def cpu_baseline_sample(frames):
    """cpu_baseline of our arm: ONE frame through the single-threaded CPU path (about 5-20 s)."""
    from oracle import oracle as O
    O.build(ref=None)
    use_ref = cpu_reference_available()
    if use_ref:
        _cpu_init(True)
    dt, _ = _cpu_worker((frames[0], use_ref))
    return {"value": 1.0 / dt, "unit": UNIT, "cores": 1, "kind": "reference" if use_ref else "port",
            "sample": "1 frame of the same workload, single thread (the reference is single-threaded); "
                      "sphere construction excluded"}


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    from pointcloud_segmentation_b200 import capi, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libh3d.so has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    frames = make_frames(FRAMES_PER_STEP, rank)
    n_pts, stride = frames.shape[1], frames.shape[2]
    pinned = torch.from_numpy(frames).pin_memory()
    on_dev = pinned.cuda()
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device="cuda")
    pipe = capi.Pipeline(GRANULARITY, synth.WINDOW, synth.LEAF, synth.OPT_DX, MINVOTES, NLINES, max_raw_points=n_pts)
    lib = capi.lib()
    lib.h3d_launch_count.restype = __import__("ctypes").c_ulonglong
    frame_bytes = n_pts * stride * 4

    def one_step(e2e):
        """one batch of frames; returns (sum of device times us, [per-phase sums], d2h bytes, lines, votes, n_filtered)"""
        flush.fill_(rank + 1)              # L2 flush between timed steps (inputs are far smaller than L2)
        torch.cuda.synchronize()
        tot, phases, d2h, lines, votes, nf = 0.0, np.zeros(4), 0, 0, 0, 0
        for k in range(FRAMES_PER_STEP):
            if e2e:
                res = pipe.run_ptr(pinned[k].data_ptr(), n_pts, stride)
            else:
                res = pipe.run_dev(on_dev[k].data_ptr(), n_pts, stride)
            us = pipe.timing_us()
            tot += us[3]
            phases += np.array(us)
            d2h += 216 + min(res.n_iters, pipe.max_iters) * 200 + 4 * res.pool_used   # state + records + inlier lists
            lines += res.n_lines
            votes += res.votes_cast
            nf += res.n_input
        return tot, phases, d2h, lines, votes, nf

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(e2e):
        for _ in range(args.warmup):
            one_step(e2e)
        barrier()
        launches0 = lib.h3d_launch_count()
        wall0 = time.perf_counter()
        acc = [one_step(e2e) for _ in range(args.steps)]
        barrier()
        wall = time.perf_counter() - wall0
        launches = lib.h3d_launch_count() - launches0
        dev_us = sum(a[0] for a in acc)
        t = torch.tensor([dev_us], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return {"dev_us": float(t.item()), "wall_s": wall, "launches": int(launches),
                "phases": sum(a[1] for a in acc), "d2h": sum(a[2] for a in acc) / args.steps,
                "lines": sum(a[3] for a in acc), "votes": sum(a[4] for a in acc), "nf": sum(a[5] for a in acc)}

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dev = timed(False)
    e2e = timed(True)
    clocks = sampler.stop() if rank == 0 else None

    n_frames_rank = FRAMES_PER_STEP * args.steps
    total_frames = n_frames_rank * world
    value = total_frames / (dev["dev_us"] * 1e-6)
    e2e_value = total_frames / (e2e["dev_us"] * 1e-6)

    # roofline of the dominant kernel (K1, the first vote = Hough::add), measured live
    from pointcloud_segmentation_b200.capi import device_info
    num_b = 20481
    n_f = dev["nf"] / n_frames_rank
    vote_us = dev["phases"][1] / n_frames_rank
    algo_bytes = 8.0 * n_f * num_b + 24.0 * n_f + 40.0 * num_b      # 8 B/vote + 24 B/point + 40 B/direction
    peak, peak_src = measured_peak()
    achieved = algo_bytes / (vote_us * 1e-6) / 1e9
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "vote_kernel_traffic.json")))["dram_bytes_per_launch"]
    except Exception:
        pass

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev["dev_us"] / 1e3 / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "configs[1]: single 224x172 ToF frames (38528 px) -> PassThrough+VoxelGrid -> "
                                   "Hough g6 (20481 directions), dx=sqrt(3)*leaf, opt_nlines 20, opt_minvotes 12",
                       "frames_per_step": FRAMES_PER_STEP, "filtered_points_per_frame": round(n_f, 1),
                       "lines_per_frame": dev["lines"] / n_frames_rank, "l2": "flushed between steps (256 MiB write)",
                       "parallelism": f"frame-parallel x{world}" if world > 1 else "single GPU",
                       "sm_count": device_info()["sm_count"]},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": frame_bytes * FRAMES_PER_STEP,
                    "d2h_bytes_per_step": int(e2e["d2h"]), "ms_per_frame": e2e["dev_us"] / 1e3 / n_frames_rank,
                    "wall_ms_per_frame": 1e3 * e2e["wall_s"] / n_frames_rank},
            "gpu_launches": dev["launches"] + e2e["launches"],
            "clocks": clocks,
            "roofline": {"kernel": "vote_resident_kernel (K1, Hough::add)", "bound": "hbm", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": algo_bytes,
                         "us_per_launch": vote_us, "gvotes_per_s": n_f * num_b / vote_us / 1e3},
            "votes_per_s": dev["votes"] / (dev["dev_us"] * 1e-6) * world,
            "phase_us_per_frame": {k: float(v) / n_frames_rank for k, v in
                                   zip(("prefilter", "first_vote", "loop", "total"), dev["phases"])},
            "wall_ms_per_frame": 1e3 * dev["wall_s"] / n_frames_rank,
        }
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline_sample(frames)
        print(json.dumps(out), flush=True)
    pipe.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the ~10 s CPU sample (profiling runs)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_cpu_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
