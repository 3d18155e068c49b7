#!/usr/bin/env python
"""bench.py -- frames/s of the iterative 3D Hough hot path on BASELINE.json config[1]:
single ToF frames 224x172 (38 528 px) -> PassThrough + VoxelGrid (leaf 0.05/1.5) -> Hough at
granularity 6, opt_nlines 20, opt_minvotes 12, dx = sqrt(3)*leaf.

  python bench.py --gpus N --steps K --warmup W            our arm (libh3d.so, CUDA)
  python bench.py --impl reference --gpus N --steps K ...  CPU arm: the reference's own code
                                                           (oracle/_ref, else the oracle port)

A step is one pass of the hot path over a batch of FRAMES_PER_STEP synthetic frames (DISTINCT_FRAMES different ones, cycled)
(poses along the tower trajectory).  Frames are independent (one Hough per frame) and the loop of
one frame is latency-bound, so a GPU keeps K frames in flight: K pipelines driven by a few host
threads of the library (h3d_replay_run, split-phase workers: --replay-threads), each limited to a share of the SMs for its persistent loop
kernels (--concurrency 24 x --sm-share 9 by default = 216 > 148 SMs, oversubscribed on purpose: a cooperative
kernel then waits for free SMs while other pipelines' kernels fill the gaps).
`value` is timed with the frames resident in HBM (h3d_pipeline_run_dev), `e2e` through
h3d_pipeline_run with pinned HOST frames: H2D of the raw frame and D2H of the line records and
inlier lists are inside the timed region.  A step is bracketed by device synchronisation and a pair
of CUDA events (all pipeline streams drained on both sides); for N > 1 every rank replays the same
frames in the same order (frame-parallel, no collective on the data path) and the maximum over ranks is
reported.  Every frame's complete result is hashed in both timed legs and compared with the same frame run alone
(`parity_in_timed_region`; a mismatch makes the run fail), and a `direction_sharded` leg runs configs[3]
(1 M points, accumulator sharded by direction over the ranks, one 8-byte NCCL all-reduce per iteration).
A third leg runs one frame at a time on all SMs: `single_frame_latency` (the < 1 ms target of
BASELINE.json's north_star) and the `roofline` of the frame's one kernel, timed alone.
Only the cpu_baseline leg and --impl reference execute anything under oracle/.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# one hardware queue per pipeline stream (default 8: streams would alias and serialise)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FRAMES_PER_STEP = int(os.environ.get("H3D_FRAMES_PER_STEP", "256"))   # frames per step: 16 passes over the DISTINCT_FRAMES frames.
                              # Every step ends by draining the pipelines (the last frames run with idle SMs beside
                              # them); with 64 frames = 4 per pipeline that tail was worth 13 % of the step and
                              # depended on the ORDER of the frames (measured: the same frames rotated by 5 on a
                              # second rank ran 13 % slower than the heaviest-first order of rank 0)
DISTINCT_FRAMES = 16
FRAMES_IN_FLIGHT = 24          # pipelines per GPU ...
SM_SHARE = 9                   # ... each limited to this many SMs for its vote and persistent loop kernels
                               # (24 x 9 = 216 > 148 SMs, oversubscribed on purpose: a cooperative loop kernel waits for
                               # 9 free SMs while the other pipelines' K0 / vote kernels and drain tails fill the gaps;
                               # sweep on B200, profiles/r02_concurrency_sweep_final.txt: 16x9 (+4 spare SMs) 4056,
                               # 20x9 4108, 24x9 4109, 28x9 4116, 16x10 4001, 18x8 3750 frames/s)
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
        try:
            f = np.load(cache)
            if f.shape == (count, synth.TOF_W * synth.TOF_H, 4):
                return f
        except Exception:   # another rank may be writing it (every rank uses the same frames): regenerate
            pass
    poses = synth.tower_trajectory(1000)
    pick = np.linspace(0, 999, count).astype(int) if rank == 0 else (np.arange(count) * 37 + 101 * rank) % 1000
    frames = np.stack([synth.tof_frame(pose=tuple(poses[k]), seed=1000 + int(k)) for k in pick]).astype(np.float32)
    try:                    # atomic: a reader never sees a half-written file
        tmp = f"{cache}.{os.getpid()}.tmp.npy"
        np.save(tmp, frames)
        os.replace(tmp, cache)
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


def check_against_golden(frames, lines):
    """the bench frames are the ones tests/golden/reference_golden_g56.npz holds reference traces of (recorded
    from oracle/_ref by tests/golden/make_golden_g56.py): compare the per-frame line counts when the raw frames
    regenerated here are bit-identical to the recorded ones."""
    path = os.path.join(ROOT, "tests", "golden", "reference_golden_g56.npz")
    try:
        gold = np.load(path)
        same = all(hashlib.sha256(np.ascontiguousarray(frames[k]).tobytes()).hexdigest() == str(gold[f"tof{k}_raw_sha256"])
                   for k in range(len(lines)))
        if not same:
            return {"frames_identical_to_golden": False}
        bad = sum(int(gold[f"tof{k}_n_lines"]) != lines[k] for k in range(len(lines)))
        return {"frames_identical_to_golden": True, "frames": len(lines), "line_count_mismatches": bad,
                "reference_lines_total": int(gold["tof_total_lines"]), "lines_total": int(sum(lines))}
    except Exception as exc:   # the golden file is test infrastructure: its absence must not stop a bench
        return {"error": str(exc)}


C4_POINTS, C4_NLINES = 1_000_000, 50
C4_DIGEST_N1 = "2a671c5540398251"   # the 50 lines of configs[3] on ONE GPU (profiles/r01_c4_direction_sharded.json n1)


def direction_sharded_leg(rank, world, dev, reps=2):
    """BASELINE.json configs[3]: the 1 M-point infrastructure cloud (200 beams), granularity 6, dx = d/64, 50 lines,
    the accumulator sharded by DIRECTION over the ranks (h3d_hough3dlines_sharded: every rank holds all points
    and directions [j0, j1); Hough::getLine is one in-stream 8-byte ncclAllReduce(MAX) per iteration over the
    torch-initialised NCCL world).  STRONG scaling: the work is fixed, N = 1 runs the same cloud on one GPU.
    Device times (CUDA events on the loop's stream), max over ranks; the digest of the detected lines is
    compared on every rank with the committed one-GPU digest."""
    import torch
    import torch.distributed as dist
    from pointcloud_segmentation_b200 import capi, sharding, synth

    def max_over_ranks(v):
        if world == 1:
            return float(v)
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    pts = synth.infrastructure(C4_POINTS)
    num_b = 20481
    capi.init_space(GRANULARITY)
    # Hough::add of the rank's shard alone (the vote kernel), CUDA events inside the library
    x, _ = capi.shift_to_origin(pts)
    mn, mx = capi.min_max(x)
    dx = float(np.sqrt(((mx - mn) ** 2).sum())) / 64.0
    j0, j1 = sharding.direction_shards(num_b, world)[rank]
    hg = capi.Hough(mn, mx, dx, shard=(j0, j1))
    hg.add(x)
    add_us = max_over_ranks(hg.bench_vote(reps=3))
    hg.close()
    del x
    ar_us, comm = 0.0, None
    if world > 1:
        t = torch.zeros(1, dtype=torch.int64, device=dev)
        for _ in range(20):
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(200):
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e1.record()
        torch.cuda.synchronize()
        ar_us = max_over_ranks(e0.elapsed_time(e1) * 1e3 / 200)
        uid = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            uid = torch.tensor(list(capi.Comm.unique_id()), dtype=torch.uint8, device=dev)
        dist.broadcast(uid, 0)
        comm = capi.Comm(bytes(uid.cpu().tolist()), rank, world)
    loop_ms, phases, wall_ms = [], None, []
    for rep in range(reps + 1):                      # the first run warms NCCL and the allocations up
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        tr = capi.hough3dlines(pts, GRANULARITY, 0.0, 10, C4_NLINES, mode=1, comm=comm)
        wall = (time.perf_counter() - t0) * 1e3
        us = capi.hough3dlines_timing()
        if rep:
            loop_ms.append(max_over_ranks(us[3] / 1e3)); wall_ms.append(max_over_ranks(wall))
            phases = [max_over_ranks(u / 1e3) for u in us]
    if comm is not None:
        comm.close()
    dig = hashlib.sha256()
    n_lines = 0
    for it in tr.iters:
        if it["status"] == 0:
            n_lines += 1
            dig.update(np.array([it["votes"], it["argmax"], len(it["inliers2"])], dtype=np.uint64).tobytes() +
                       np.asarray(it["a2"]).tobytes() + np.asarray(it["b2"]).tobytes())
    digest = dig.hexdigest()[:16]
    ok = torch.tensor([1.0 if digest == C4_DIGEST_N1 else 0.0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    votes = C4_POINTS * num_b
    return {
        "workload": "configs[3]: 1M-point infrastructure cloud (200 beams), g6, dx=d/64, opt_nlines 50, accumulator "
                    "sharded by direction over the ranks, all points on every rank",
        "scaling": "strong", "n_gpus": world, "directions_per_rank": j1 - j0,
        "add_ms": add_us / 1e3, "add_gvotes_per_s": votes / add_us / 1e3,
        "add_roofline_frac": (8.0 * votes + 24.0 * C4_POINTS + 40.0 * num_b) / (add_us * 1e-6) / 1e9 / (measured_peak()[0] * world),
        "allreduce_us": ar_us, "loop_ms": float(np.min(loop_ms)), "loop_ms_runs": loop_ms,
        "loop_wall_ms": float(np.min(wall_ms)),
        "loop_phase_ms": dict(zip(("upload_prep", "first_vote", "loop_kernels", "total"), phases)),
        "lines": n_lines, "lines_digest": digest, "digest_n1_golden": C4_DIGEST_N1,
        "digest_matches_n1": bool(ok.item() == 1.0),
        "timed": "CUDA events on the loop's stream around the whole h3d_hough3dlines_sharded call (host points in, "
                 "records out), max over ranks, best of %d runs after one warm-up" % reps,
    }


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
    if use_ref:
        eng.set_index_recovery(False)      # time the reference calls only: no flat-index scan, no index lists
    t0 = time.perf_counter()
    filt = orc.prefilter(frame, synth.WINDOW, float(synth.LEAF))
    t1 = time.perf_counter()
    tr = eng.hough3dlines(filt.astype(np.float64), GRANULARITY, synth.OPT_DX, MINVOTES, NLINES, mode=0)
    t2 = time.perf_counter()
    phases = np.concatenate([[t1 - t0], np.asarray(tr.timing_us) * 1e-6]) if tr.timing_us is not None else np.zeros(7)
    return t2 - t0, tr.n_lines, phases


CPU_PHASES = ("prefilter", "add", "subtract", "getLine", "pointsCloseToLine", "orthogonal_LSQ", "removePoints")


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
    frames = make_frames(DISTINCT_FRAMES)
    with mp.get_context("spawn").Pool(workers, initializer=_cpu_init, initargs=(use_ref,)) as pool:
        jobs = [(frames[k % len(frames)], use_ref) for k in range(workers)]
        for _ in range(args.warmup):
            pool.map(_cpu_worker, jobs)
        t0 = time.perf_counter()
        lines = 0
        phases, busy = np.zeros(len(CPU_PHASES)), 0.0
        for _ in range(args.steps):
            for r in pool.map(_cpu_worker, jobs):
                lines += r[1]; phases += r[2]; busy += r[0]
        dt = time.perf_counter() - t0
    value = workers * args.steps / dt
    n_done = workers * args.steps
    phase_ms = {k: 1e3 * float(v) / n_done for k, v in zip(CPU_PHASES, phases)}
    phase_ms["other (Hough ctor zero-fill, copies into PointCloud, driver)"] = 1e3 * (busy - float(phases.sum())) / n_done
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "configs[1]: single 224x172 ToF frames (38528 px) -> PassThrough+VoxelGrid -> "
                               "Hough g6 (20481 directions), dx=sqrt(3)*leaf, opt_nlines 20, opt_minvotes 12",
                   "frames_per_step": workers, "parallelism": f"{workers} single-threaded processes, one per host core",
                   "lines_per_frame": lines / max(1, workers * args.steps)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": "reference" if use_ref else "port",
                         "sample": f"{workers} frames per step, one single-threaded process per core",
                         "ms_per_frame_by_phase": phase_ms,
                         "timed": "reference calls only (add/subtract/getLine/pointsCloseToLine/removePoints of the "
                                  "unmodified library; orthogonal_LSQ and the pre-filter are the stated restatements); "
                                  "the shim's flat-index recovery scan and index lists are switched off"},
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
    dt, _, phases = _cpu_worker((frames[0], use_ref))
    return {"value": 1.0 / dt, "unit": UNIT, "cores": 1, "kind": "reference" if use_ref else "port",
            "ms_by_phase": {k: 1e3 * float(v) for k, v in zip(CPU_PHASES, phases)},
            "sample": "1 frame of the same workload, single thread (the reference is single-threaded); "
                      "sphere construction excluded"}


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def run_gpu_arm(args):
    import ctypes
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
        if os.environ.get("H3D_BENCH_BACKEND") == "gloo":      # experiment switch (tools/): plumbing without NCCL
            dist.init_process_group("gloo")
        elif os.environ.get("H3D_BENCH_BACKEND") == "nccl-lazy":
            dist.init_process_group("nccl")
        else:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    # every rank replays the SAME 16 frames in the same order (weak scaling: per-GPU work fixed as N grows; with a
    # frame set or an order of its own per rank the slowest rank's drain tail set the step time: 0.89 "efficiency"
    # at N = 2 that was data, not the machine -- profiles/r02_n2_efficiency_experiments.txt)
    frames = make_frames(DISTINCT_FRAMES, 0)
    rot = 0
    n_pts, stride = frames.shape[1], frames.shape[2]
    pinned = torch.from_numpy(frames).pin_memory()
    on_dev = pinned.cuda()
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device="cuda")
    lib = capi.lib()
    lib.h3d_launch_count.restype = ctypes.c_ulonglong
    sm_count = capi.device_info()["sm_count"]
    frame_bytes = n_pts * stride * 4
    K = max(1, min(args.concurrency, FRAMES_PER_STEP))
    # host threads: K per rank.  They spin on their streams while the box has a core for each of them;
    # with several ranks per box they poll and yield the core between polls (h3d_set_wait_mode 2;
    # measured on a 32-core 8-GPU box: 8 GPUs 11.7 k frames/s spinning, 22.2 k blocking, 22.9 k yielding)
    try:
        host_cores = len(os.sched_getaffinity(0))
    except AttributeError:
        host_cores = os.cpu_count() or 1
    # auto = poll + yield whenever frames are in flight: equal to spinning while cores are free
    # (3.31 k vs 3.31 k frames/s at 1 GPU / 16 cores), far ahead of it when they are not
    wait_mode = {"spin": 0, "block": 1, "yield": 2}.get(args.wait, 2 if K > 1 else 0)
    capi.set_wait_mode(wait_mode)
    # fewer cores than ranks x K: each host thread drives several pipelines in split phase (submit, poll, complete)
    cores_per_rank = max(1, host_cores // world)
    # (measured at N = 1 on 16 cores: 16 / 8 / 4 / 2 threads for 16 pipelines -> 3763 / 3783 / 3784 / 3772 frames/s:
    # a frame costs the host ~30 us of launches and one copy, so a few threads are enough on any box)
    replay_threads = args.replay_threads if args.replay_threads >= 0 else min(K, max(2, min(4, cores_per_rank - 1)))
    capi.set_replay_threads(replay_threads)
    share = 0 if K == 1 else (args.sm_share if args.sm_share > 0 else max(1, sm_count // K))

    def new_pipe(sm_share):
        p = capi.Pipeline(GRANULARITY, synth.WINDOW, synth.LEAF, synth.OPT_DX, MINVOTES, NLINES, max_raw_points=n_pts)
        p.set_sm_share(sm_share)
        return p

    # K pipelines, K host threads inside the library (h3d_replay_run), `share` SMs each for the
    # persistent loop kernel: frames are independent (one Hough per frame, wrap.h:220/347) and a
    # frame's loop is latency-bound, so the GPU is filled with frames in flight
    replay = capi.Replay(K, share, GRANULARITY, synth.WINDOW, synth.LEAF, synth.OPT_DX, MINVOTES, NLINES, n_pts,
                         FRAMES_PER_STEP)
    # SMs left over by K x share (148 - 16 x 9 = 4) go to the first pipelines, one each
    spare = sm_count - K * share if (K > 1 and share > 0 and not args.no_spare_sms) else 0
    for k in range(max(0, min(spare, K))):
        replay.pipes[k].set_sm_share(share + 1)
    host_ptrs = [pinned[(k + rot) % DISTINCT_FRAMES].data_ptr() for k in range(FRAMES_PER_STEP)]
    dev_ptrs = [on_dev[(k + rot) % DISTINCT_FRAMES].data_ptr() for k in range(FRAMES_PER_STEP)]
    counts = [n_pts] * FRAMES_PER_STEP
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    iter_bytes = ctypes.sizeof(capi.Iter)

    def frame_digest(res, iters_base, pool_arr, pool_base, max_iters):
        """sha256 over everything a frame returns: header counts, every iteration record (votes, argmax,
        status, fitted lines, inlier counts/offsets) and both inlier index lists of every iteration"""
        n_rec = min(res.n_iters, max_iters)
        h = hashlib.sha256()
        h.update(np.array([res.rc, res.n_iters, res.n_lines, res.n_input, res.num_x, res.pool_used], dtype=np.int64).tobytes())
        h.update(ctypes.string_at(iters_base, n_rec * iter_bytes))
        h.update(pool_arr[pool_base:pool_base + int(res.pool_used)].tobytes())
        return h.digest()

    # what every frame must return: each distinct frame once through ONE pipeline on all SMs, before any timing
    want_digest, want_lines = [], []
    ref_pipe = new_pipe(0)
    for k in range(DISTINCT_FRAMES):
        r = ref_pipe.run_dev(on_dev[k].data_ptr(), n_pts, stride)
        want_digest.append(frame_digest(r, ctypes.addressof(ref_pipe.iters), ref_pipe.pool, 0, ref_pipe.max_iters))
        want_lines.append(int(r.n_lines))
    ref_pipe.close()
    golden = check_against_golden(frames, want_lines)   # every rank holds the golden frames

    def one_step(e2e):
        """one batch of FRAMES_PER_STEP frames over the K pipelines; device time of the whole batch (ms)"""
        flush.fill_(rank + 1)              # L2 flush between timed steps (inputs are far smaller than L2)
        torch.cuda.synchronize()
        ev0.record()
        res = replay.run(host_ptrs if e2e else dev_ptrs, counts, stride, on_device=not e2e)
        torch.cuda.synchronize()           # every pipeline stream has drained
        ev1.record()
        ev1.synchronize()
        out = dict(d2h=0, lines=0, votes=0, nf=0, iters=0, mismatches=0, checked=0)
        base = ctypes.addressof(replay.iters)
        for k in range(FRAMES_PER_STEP):
            r = res[k]
            n_rec = min(r.n_iters, replay.max_iters)
            out["d2h"] += 216 + n_rec * 200 + 4 * r.pool_used          # state + records + inlier lists
            out["lines"] += r.n_lines; out["votes"] += r.votes_cast; out["nf"] += r.n_input; out["iters"] += r.n_iters
            # parity inside the timed legs: the frame's complete result against the one-pipeline run of the same frame
            got = frame_digest(r, base + k * replay.max_iters * iter_bytes, replay.pool, k * replay.pool_cap, replay.max_iters)
            out["checked"] += 1
            out["mismatches"] += got != want_digest[(k + rot) % DISTINCT_FRAMES]
        out["ms"] = ev0.elapsed_time(ev1)
        return out

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
        acc = [one_step(e2e) for _ in range(args.steps)]
        barrier()
        launches = lib.h3d_launch_count() - launches0
        t = torch.tensor([sum(a["ms"] for a in acc)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out = {k: sum(a[k] for a in acc) for k in acc[0]}
        out["ms"] = float(t.item())
        out["launches"] = int(launches)
        return out

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dev = timed(False)
    e2e = timed(True)
    clocks = sampler.stop() if rank == 0 else None
    replay.close()
    capi.set_wait_mode(0)                  # one frame at a time: lowest latency
    # ---- single-frame latency + roofline leg: ONE pipeline with all SMs, frames resident in HBM ---------
    num_b = 20481
    lat = new_pipe(0)
    lat_us, phase = [], np.zeros(4)
    algo_bytes = algo_dense = k1_bytes_sum = 0.0
    votes_add = votes_sub = 0
    k1_best = None
    for rep in range(2 + max(1, args.steps // 2)):
        flush.fill_(7)
        torch.cuda.synchronize()
        for k in range(DISTINCT_FRAMES):
            res = lat.run_dev(on_dev[k].data_ptr(), n_pts, stride)
            if rep < 2:
                continue                                         # warm-up
            us = lat.timing_us()
            lat_us.append(us[3]); phase += np.array(us)
            n_alive, cells = res.n_input, res.num_x * res.num_x * num_b
            k1b = 8.0 * n_alive * num_b + 24.0 * n_alive + 40.0 * num_b        # K1: Hough::add
            k1_bytes_sum += k1b
            cand = {"us": float(us[1]), "points": int(n_alive), "gvotes_per_s": n_alive * num_b / us[1] / 1e3,
                    "achieved_gbs": k1b / us[1] / 1e3}
            if k1_best is None or cand["achieved_gbs"] > k1_best["achieved_gbs"]:
                k1_best = cand
            votes_add += n_alive * num_b
            b = 24.0 * n_alive + 9.0 * num_b                                   # loop prologue: points + records
            dense = 0.0
            for q in range(min(res.n_iters, lat.max_iters)):
                r = lat.iters[q]
                b += 9.0 * num_b + 25.0 * n_alive * (1 if r.status == 1 else 2)   # K2 (records) + K3 passes
                dense += 4.0 * cells                                              # K2 as the reference scans it
                if r.status == 0:
                    b += 8.0 * r.n2 * num_b + 24.0 * r.n2 + 40.0 * num_b          # K4: Hough::subtract
                    votes_sub += r.n2 * num_b
                    n_alive -= r.n2
            algo_bytes += b; algo_dense += b + dense
    n_lat = len(lat_us)
    # north_star: "a single-GPU 38k-point ToF frame ... processed end-to-end in under 1 ms".  End to end = the
    # C-ABI call a consumer makes (h3d_pipeline_run): pinned HOST frame in, line records and inlier lists in HOST
    # memory out; host wall clock around the call, so H2D, every launch, the wait and the copies are inside.
    e2e_lat = []
    for rep in range(2 + max(1, args.steps // 2)):
        flush.fill_(9)
        torch.cuda.synchronize()
        for k in range(DISTINCT_FRAMES):
            t0 = time.perf_counter()
            lat.run_ptr(pinned[k].data_ptr(), n_pts, stride)
            dt = time.perf_counter() - t0
            if rep >= 2:
                e2e_lat.append(dt * 1e3)
    # the same frames with the segment post-tests (wrap.h:261-337) switched on
    seg_leg = None
    try:
        lat.enable_segments(capi.SegmentConfig.make(float(np.sqrt(3.0)) * float(synth.LEAF), 0.995, 1.5, [0.05]))
        seg_us, tested, accepted = [], 0, 0
        for rep in range(3):
            for k in range(DISTINCT_FRAMES):
                lat.run_dev(on_dev[k].data_ptr(), n_pts, stride)
                if rep == 0:
                    continue
                seg_us.append(lat.timing_us()[3])
                _, segs = lat.segments()
                tested += len(segs); accepted += sum(1 for s_ in segs if s_.status == 0)
        lat.enable_segments(None)
        seg_leg = {"ms_mean": float(np.mean(seg_us)) / 1e3, "extra_us_per_frame": float(np.mean(seg_us) - np.mean(lat_us)),
                   "lines_tested_per_frame": tested / len(seg_us), "segments_kept_per_frame": accepted / len(seg_us),
                   "note": "single-frame latency leg with h3d_pipeline_enable_segments: one extra launch per frame"}
    except Exception as exc:
        seg_leg = {"error": str(exc)}
    # K1 (Hough::add) is its own launch: CUDA events around it (h3d_pipeline_timing [1])
    k1_us_mean = phase[1] / n_lat
    k1 = {"kernel": "vote_resident_kernel (one launch per frame)", "us_per_launch": float(k1_us_mean),
          "algorithmic_bytes_per_launch": k1_bytes_sum / n_lat,
          "achieved": k1_bytes_sum / n_lat / k1_us_mean / 1e3, "unit": "GB/s",
          "gvotes_per_s": votes_add / n_lat / k1_us_mean / 1e3, "best_frame": k1_best}
    k1["frac"] = k1["achieved"] / measured_peak()[0]
    lat.close()

    # ---- configs[3]: the direction-sharded loop over the same NCCL world (every N, N = 1 included) ---------
    sharded = None
    if not args.no_direction_sharded:
        try:
            sharded = direction_sharded_leg(rank, world, torch.device("cuda", local))
        except Exception as exc:   # reported, and fatal below: a missing leg must not pass silently
            sharded = {"error": f"{type(exc).__name__}: {exc}"}
    n_frames_rank = FRAMES_PER_STEP * args.steps
    total_frames = n_frames_rank * world
    golden_bad = int(golden.get("line_count_mismatches", 0)) if isinstance(golden, dict) else 0   # this rank's GPU vs the reference's line counts
    parity = torch.tensor([dev["checked"] + e2e["checked"], dev["mismatches"] + e2e["mismatches"] + golden_bad],
                          dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(parity, op=dist.ReduceOp.SUM)
    parity = parity.tolist()
    value = total_frames / (dev["ms"] * 1e-3)
    e2e_value = total_frames / (e2e["ms"] * 1e-3)
    peak, peak_src = measured_peak()
    kernel_us = phase[2] / n_lat
    achieved = algo_bytes / n_lat / (kernel_us * 1e-6) / 1e9
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "fused_kernel_traffic.json")))["dram_bytes_per_launch"]
    except Exception:
        pass

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev["ms"] / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "configs[1]: single 224x172 ToF frames (38528 px) -> PassThrough+VoxelGrid -> "
                                   "Hough g6 (20481 directions), dx=sqrt(3)*leaf, opt_nlines 20, opt_minvotes 12",
                       "frames_per_step": FRAMES_PER_STEP, "filtered_points_per_frame": round(dev["nf"] / n_frames_rank, 1),
                       "lines_per_frame": dev["lines"] / n_frames_rank, "l2": "flushed between steps (256 MiB write)",
                       "distinct_frames": DISTINCT_FRAMES, "frames_in_flight_per_gpu": K, "sms_per_frame": share or sm_count, "pipelines_with_one_more_sm": max(0, min(spare, K)),
                       "parallelism": (f"frame-parallel x{world} GPUs, " if world > 1 else "single GPU, ") +
                                      f"{K} frames in flight per GPU",
                       "sm_count": sm_count, "host_cores": host_cores,
                       "host_wait": ("spin", "blocking events", "poll + yield")[wait_mode],
                       "host_threads_per_rank": replay_threads or K},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": frame_bytes * FRAMES_PER_STEP,
                    "d2h_bytes_per_step": int(e2e["d2h"] / args.steps), "ms_per_step": e2e["ms"] / args.steps,
                    "lines_per_frame": e2e["lines"] / n_frames_rank},
            "parity_in_timed_region": {
                "frames": int(parity[0]), "mismatches": int(parity[1]),
                "checked": "sha256 of every frame's complete result (iteration records + both inlier lists) in BOTH "
                           "timed legs, all ranks, against the same frame run alone through one pipeline",
                "expected_lines_per_frame": sum(want_lines) / DISTINCT_FRAMES, "golden": golden},
            "gpu_launches": dev["launches"] + e2e["launches"],
            "clocks": clocks,
            "single_frame_latency": {"ms_mean": float(np.mean(lat_us)) / 1e3, "ms_max": float(np.max(lat_us)) / 1e3,
                                     "frames": n_lat, "note": "one frame at a time on all SMs, frame resident in HBM",
                                     "e2e_ms_mean": float(np.mean(e2e_lat)), "e2e_ms_max": float(np.max(e2e_lat)),
                                     "e2e_ms_p50": float(np.median(e2e_lat)),
                                     "e2e_note": "h3d_pipeline_run: pinned host frame in (616 KB H2D), records + inlier lists "
                                                 "in host memory out; host wall clock around the call, L2 flushed per pass",
                                     "phase_us": {k: float(v) / n_lat for k, v in
                                                  zip(("prefilter", "first_vote_kernel", "fused_loop_kernel", "total"), phase)}},
            "roofline": {"kernel": "fused_loop_kernel (all iterations: K2 argmax + K3 fit + K4 un-vote/re-vote, one launch per frame)",
                         "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": algo_bytes / n_lat, "us_per_launch": kernel_us,
                         "timed": "CUDA events around the launch, one frame at a time (single_frame_latency leg)",
                         "note": "SURVEY 8(d) bytes: 8 B/vote (subtract) + 24 B/point + 40 B/direction, "
                                 "25 B/live point per K3 pass, K2 counted on the 9 B/direction records",
                         "achieved_dense_equivalent": algo_dense / n_lat / (kernel_us * 1e-6) / 1e9,
                         "k1_first_vote_kernel": k1},
            "direction_sharded": sharded,
            "segment_post_tests": seg_leg,
            "votes_per_s": (votes_add + votes_sub) / n_lat / ((kernel_us + k1_us_mean) * 1e-6),
            "votes_evaluated_per_s_throughput": dev["votes"] / (dev["ms"] * 1e-3) * world,
        }
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline_sample(frames)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if sharded is not None and (sharded.get("error") or not sharded.get("digest_matches_n1")):
        raise SystemExit(f"bench.py: direction-sharded leg failed on rank {rank}: {sharded}")
    if parity[1] > 0:
        raise SystemExit(f"bench.py: {int(parity[1])} of {int(parity[0])} frames in the timed legs differ from the "
                         "one-pipeline result of the same frame")


# ---------------------------------------------------------------------------------------------
# --workload c3: BASELINE.json configs[2] as written
# ---------------------------------------------------------------------------------------------
def _c3_frame(k):
    from pointcloud_segmentation_b200 import synth
    poses = synth.tower_trajectory(1000)
    return synth.tof_frame(pose=tuple(poses[k]), seed=1000 + int(k)).astype(np.float32)


def _c3_oracle(frame):
    from oracle import oracle as O
    from pointcloud_segmentation_b200 import synth
    orc = O.Oracle()
    filt = orc.prefilter(frame, synth.WINDOW, float(synth.LEAF))
    tr = orc.hough3dlines(filt.astype(np.float64), GRANULARITY, synth.OPT_DX, MINVOTES, NLINES, mode=0)
    return [(int(it.votes), int(it.argmax), int(it.status), np.asarray(it.inliers1, dtype=np.uint32).tobytes(),
             np.asarray(it.inliers2, dtype=np.uint32).tobytes(), np.asarray(it.a2).tobytes(), np.asarray(it.b2).tobytes())
            for it in tr.iters]


def run_c3(args):
    """configs[2]: batched replay of 1000 DISTINCT synthetic ToF frames along the tower trajectory, once, sharded
    frame-parallel over the N GPUs (frame i -> rank i mod N, no collective on the data path).  Host frames in,
    line records + inlier lists out (the e2e form); device time = CUDA events around every h3d_replay_run call,
    summed, max over ranks.  Parity: 50 frames of rank 0 (every 20th of the trajectory) are recomputed by the CPU
    oracle after the timed region and compared record for record."""
    import ctypes
    import multiprocessing as mp
    import torch
    import torch.distributed as dist
    from pointcloud_segmentation_b200 import capi, sharding, synth

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libh3d.so has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    total = 1000
    mine = sharding.frames_for_rank(total, rank, world)
    with mp.get_context("spawn").Pool(max(1, min(16, (os.cpu_count() or 1) // world))) as pool:
        frames = np.stack(pool.map(_c3_frame, mine, chunksize=8))
    n_pts, stride = frames.shape[1], frames.shape[2]
    pinned = torch.from_numpy(frames).pin_memory()
    K, chunk = args.concurrency, 250   # frames per timed call: the drain tail of a call is one frame time
    capi.set_wait_mode(2 if K > 1 else 0)
    try:
        cores_per_rank = max(1, len(os.sched_getaffinity(0)) // world)
    except AttributeError:
        cores_per_rank = max(1, (os.cpu_count() or 1) // world)
    capi.set_replay_threads(args.replay_threads if args.replay_threads >= 0 else min(K, max(2, min(4, cores_per_rank - 1))))
    replay = capi.Replay(K, args.sm_share if K > 1 else 0, GRANULARITY, synth.WINDOW, synth.LEAF, synth.OPT_DX, MINVOTES,
                         NLINES, n_pts, chunk)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    warm = [pinned[k % len(mine)].data_ptr() for k in range(chunk)]
    for _ in range(2):
        replay.run(warm, [n_pts] * chunk, stride, on_device=False)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms, lines, got = 0.0, 0, {}
    for c0 in range(0, len(mine), chunk):
        idx = list(range(c0, min(c0 + chunk, len(mine))))
        torch.cuda.synchronize()
        ev0.record()
        res = replay.run([pinned[k].data_ptr() for k in idx], [n_pts] * len(idx), stride, on_device=False)
        torch.cuda.synchronize()
        ev1.record(); ev1.synchronize()
        ms += ev0.elapsed_time(ev1)
        for q, k in enumerate(idx):
            lines += res[q].n_lines
            if rank == 0 and mine[k] % 20 == 0:
                its = replay.iters[q * replay.max_iters:(q + 1) * replay.max_iters]
                plq = replay.pool[q * replay.pool_cap:(q + 1) * replay.pool_cap]
                tr = capi.Trace(res[q], its, plq)
                got[k] = [(it["votes"], it["argmax"], it["status"], it["inliers1"].tobytes(), it["inliers2"].tobytes(),
                           it["a2"].tobytes(), it["b2"].tobytes()) for it in tr.iters]
    replay.close()
    t = torch.tensor([ms, float(lines)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        ms = float(tmax[0].item())
    lines_total = int(t[1].item())
    if rank == 0:
        from concurrent.futures import ThreadPoolExecutor
        from oracle import oracle as O
        O.build(ref=False)
        keys = sorted(got)
        t0 = time.perf_counter()
        with ThreadPoolExecutor(max_workers=min(len(keys), os.cpu_count() or 1)) as ex:
            want = list(ex.map(_c3_oracle, [frames[k] for k in keys]))
        cpu_s = time.perf_counter() - t0
        mism = sum(1 for k, w in zip(keys, want) if got[k] != w)
        out = {"metric": "frames/s (configs[2]: replay of 1000 distinct ToF frames, frame-parallel over the GPUs)",
               "value": total / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "frames": total, "ms_total": ms,
               "higher_is_better": True, "scaling": "strong", "dtype": "f64", "data": "synthetic",
               "config": {"workload": "configs[2]: 1000 distinct 224x172 ToF frames along the tower trajectory, each "
                                      "PassThrough+VoxelGrid -> Hough g6, opt_nlines 20, opt_minvotes 12; frame i on rank i mod N",
                          "frames_in_flight_per_gpu": K, "sms_per_frame": args.sm_share if K > 1 else "all",
                          "host_frames": "pinned host memory; H2D of every frame and D2H of its records inside the timed region"},
               "lines_per_frame": lines_total / total,
               "parity_vs_oracle": {"frames": len(keys), "mismatches": mism, "oracle_cpu_s": cpu_s,
                                    "checked": "every iteration record (votes, argmax, status, both inlier index lists, a, b) "
                                               "of every 20th trajectory frame against oracle/h3d_oracle.c"}}
        print(json.dumps(out), flush=True)
        if mism:
            raise SystemExit(f"bench.py --workload c3: {mism} of {len(keys)} sampled frames differ from the oracle")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--concurrency", type=int, default=FRAMES_IN_FLIGHT,
                    help="frames in flight per GPU (K pipelines, K host threads); 1 = lowest latency")
    ap.add_argument("--wait", choices=("auto", "spin", "block", "yield"), default="auto",
                    help="how the K host threads wait for the device (auto: poll + yield when frames are in flight)")
    ap.add_argument("--no-spare-sms", action="store_true", help="do not hand the SMs left over by K x share to the first pipelines")
    ap.add_argument("--replay-threads", type=int, default=-1,
                    help="host threads per rank driving the frames in flight (-1: one per frame in flight while the "
                         "box has the cores, fewer otherwise)")
    ap.add_argument("--sm-share", type=int, default=SM_SHARE, help="SMs per frame in flight (0: sm_count / concurrency)")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the ~10 s CPU sample (profiling runs)")
    ap.add_argument("--no-direction-sharded", action="store_true",
                    help="skip the configs[3] leg (1 M points, accumulator sharded by direction over the ranks)")
    ap.add_argument("--workload", choices=("c2", "c3"), default="c2",
                    help="c2 (default): BASELINE configs[1] frames, the contract line; c3: configs[2] as written, "
                         "1000 distinct frames replayed once with an oracle parity sample (takes a few minutes)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_cpu_arm(args)
    elif args.workload == "c3":
        run_c3(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
