#!/usr/bin/env python
"""bench.py — headline benchmark of the scan-matching hot path (BASELINE.json).

Metric M1 (headline `value`): scan-to-TSDF residual + Jacobian points/s on config C2 =
"multi-resolution TSDF registration: 64-scan batch x 100k pts, high+low-res grids, 1 GPU":
one step = one Evaluate of 64 two-pose interpolated multi-resolution TSDF residual blocks
(InterpolatedMultiResolutionTSDFSpaceCostFunction3D) over 8 control poses, 6.4 M points,
residuals + all four Jacobian blocks written in Ceres' layout.
Metric M2 (reported in `correlative`): correlative candidates scored/s on config C4 (MatchFullSubmap at
pose_graph.lua's thresholds, discrete scans dealt over the ranks; the timed region includes the NCCL exchange of
the dive bound and of the best records) in two variants, rotational prefilter on ("configured") and off ("full
sweep", the ~8 M-candidate stress case the value and the roofline are quoted on); `local_search` = C3 (batched local
Match, 100 node poses at N = 2048 and 10 at N = 20480); `constraint_batch` = C5 (10 000 pairs). Selected candidates
are asserted against the CPU oracle's frozen answers (tests/golden/c4_golden.json).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

`--impl reference` times the CPU restatement of the reference (oracle/, the reference itself
cannot be compiled here — see DESIGN.md) on the host cores for the same metric and config.
"""
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

NUM_SCANS = 64
POINTS_PER_SCAN = 100000
NUM_POSES = 8
HI_RES, LO_RES = 0.10, 0.45
REL_TRUNC, MAX_WEIGHT = 2.5, 1000.0
# two-pose functor, fine level touched: 12 B point + 32 B (8 TSDFVoxel) + 120 B (r + 14 J).
ALG_BYTES_PER_POINT = 164


_JSON_OUT = None


def emit(line):
    out = _JSON_OUT if _JSON_OUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
            return int(json.load(f)[kernel]["dram_bytes_per_launch"])
    except Exception:
        return None


def build_workload(seed=2, num_scans=NUM_SCANS, points_per_scan=POINTS_PER_SCAN):
    """C2 inputs: grids as voxel lists, 64 scans in their sensor frames, control poses, block table."""
    from hectorgrapher_b200 import synth
    rng = np.random.RandomState(seed)
    sph = synth.spheres()
    map_poses = [synth.corkscrew_pose(k, NUM_POSES) for k in range(NUM_POSES)]
    hi = synth.build_tsdf(HI_RES, map_poses, 48, 1024, sph, REL_TRUNC, MAX_WEIGHT)
    lo = synth.build_tsdf(LO_RES, map_poses, 48, 1024, sph, REL_TRUNC, MAX_WEIGHT)
    clouds, i0, i1, ratios = [], [], [], []
    rings, az = 128, 800
    for s in range(num_scans):
        u = s * (NUM_POSES - 1) / num_scans
        a = int(np.floor(u))
        f = rng.uniform(0.0, 1.0)
        (oa, ya), (ob, yb) = map_poses[a], map_poses[a + 1]
        origin = oa + (ob - oa) * f
        yaw = ya + (yb - ya) * f
        pts = synth.scan(origin, yaw, rings, az, sph, rng)
        if len(pts) < points_per_scan:
            pts = np.concatenate([pts, pts[: points_per_scan - len(pts)]])
        clouds.append(pts[:points_per_scan])
        i0.append(a)
        i1.append(a + 1)
        ratios.append(f)
    poses = []
    for k in range(NUM_POSES):
        o, y = map_poses[k]
        t = o + rng.uniform(-0.05, 0.05, 3)
        q = synth.quat_mul(synth.quat_from_yaw(y), synth.small_rotation(rng, 0.02))
        poses.append(np.concatenate([t, q]))
    points = np.ascontiguousarray(np.concatenate(clouds), np.float32)
    offsets = np.arange(num_scans + 1, dtype=np.int64) * points_per_scan
    scales = np.full(num_scans, 1.0 / np.sqrt(points_per_scan))
    return dict(hi=hi, lo=lo, points=points, offsets=offsets, scales=scales, ratios=np.array(ratios),
                poses=np.array(poses), i0=np.array(i0, np.int32), i1=np.array(i1, np.int32), synth=synth,
                sph=sph, map_poses=map_poses)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.samples = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(index), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            p = [x.strip() for x in s.split(",")]
            if len(p) < 6:
                continue
            try:
                sm.append(float(p[0]))
                mx.append(float(p[1]))
            except ValueError:
                continue
            for n, v in zip(names, p[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(np.max(mx)) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class CpuBlocks:
    """The CPU restatement (oracle/) of the same functor on `threads` DIFFERENT residual blocks of the workload
    (block t on thread t: each thread walks its own 100k-point scan, like a multi-threaded Ceres evaluating the
    window's blocks, or ConstraintBuilder3D's pool)."""

    def __init__(self, w, threads):
        from oracle import pyoracle as o
        self.threads = threads
        self.hi = o.Grid.tsdf(HI_RES, REL_TRUNC, MAX_WEIGHT)
        self.hi.set_raw(*w["hi"])
        self.lo = o.Grid.tsdf(LO_RES, REL_TRUNC, MAX_WEIGHT)
        self.lo.set_raw(*w["lo"])
        self.costs, self.params = [], []
        nb = len(w["offsets"]) - 1
        for t in range(threads):
            b = t % nb
            p0, p1 = int(w["offsets"][b]), int(w["offsets"][b + 1])
            self.costs.append(o.Cost("interpolated_multires_tsdf", w["scales"][b], w["points"][p0:p1],
                                     [self.hi, self.lo], float(w["ratios"][b])))
            self.params.append(np.concatenate([w["poses"][w["i0"][b]], w["poses"][w["i1"][b]]]))
        self.n = int(w["offsets"][1])

    def run(self, reps):
        """Every thread evaluates its block `reps` times (residuals + Jacobians); returns the wall time."""
        from concurrent.futures import ThreadPoolExecutor
        t0 = time.perf_counter()
        with ThreadPoolExecutor(self.threads) as ex:
            list(ex.map(lambda k: self.costs[k].time(self.params[k], reps, 1), range(self.threads)))
        return time.perf_counter() - t0


def cpu_baseline_points(w, seconds=12.0, threads=None):
    """Times the CPU restatement on a bounded sample: `threads` different 100k-point blocks of the workload, one
    per host thread, evaluated concurrently. Returns (points/s, description, threads, CpuBlocks)."""
    threads = threads or os.cpu_count() or 1
    blocks = CpuBlocks(w, threads)
    t1 = blocks.costs[0].time(blocks.params[0], 1, 1)
    one_thread = blocks.n / t1
    tt = blocks.run(1)
    reps = max(1, int(seconds / max(tt * 1.1, 1e-3)))
    tt = blocks.run(reps)
    rate = blocks.n * reps * threads / tt
    desc = ("blocks 0..%d of the workload (100000 pts each, two-pose multi-res TSDF, residuals+Jacobians), one per "
            "thread, %d threads x %d evaluations; 1 thread: %.3g pts/s" % (threads - 1, threads, reps, one_thread))
    return rate, desc, threads, blocks


def run_reference(args, rank, world):
    if rank != 0:
        return
    w = build_workload()
    rate, desc, threads, blocks = cpu_baseline_points(w, seconds=2.0)
    n = blocks.n
    # steps: each a bounded sample of the workload (one Evaluate of `threads` different 100k-point blocks)
    for _ in range(args.warmup):
        blocks.run(1)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        blocks.run(1)
    dt = time.perf_counter() - t0
    value = n * threads * args.steps / dt
    line = {
        "impl": "reference", "metric": "scan-to-TSDF residual+Jacobian points/s", "value": value,
        "unit": "points/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C2 multi-resolution TSDF registration: 64 scans x 100k pts, 0.10 m + 0.45 m TSDF, "
                               "two-pose interpolated multi-res functor",
                   "sample_per_step": "%d of the 64 blocks (100k points each), one per host thread, 1 Evaluate each" % threads},
        "cpu_baseline": {"value": value, "unit": "points/s", "cores": threads, "kind": "port",
                         "sample": "CPU restatement (oracle/) of the reference functor; " + desc},
        "e2e": {"value": value, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


# pose_graph.lua:20, 44-45 (constraint_builder min_score, fast_correlative_scan_matcher_3d thresholds)
POSE_GRAPH_MIN_SCORE = 0.55
POSE_GRAPH_MIN_ROTATIONAL_SCORE = 0.77
POSE_GRAPH_MIN_LOW_RESOLUTION_SCORE = 0.55
HISTOGRAM_SIZE = 120  # pose_graph.lua: rotational_histogram_size


def submap_histogram(w):
    """The submap's rotational histogram the way the reference accumulates it (mapping/3d/submap_3d.cc:
    every inserted scan's histogram rotated into the submap frame and added): one coarse scan per map pose."""
    synth = w["synth"]
    total = np.zeros(HISTOGRAM_SIZE, np.float32)
    for k, (origin, yaw) in enumerate(w["map_poses"]):
        cloud = synth.scan(origin, yaw, 32, 256, w["sph"], np.random.RandomState(700 + k))
        pts = synth.voxel_downsample(cloud, 0.2, 2048, seed=8)
        total += synth.rotate_histogram(synth.rotational_histogram(pts, HISTOGRAM_SIZE), np.float32(yaw))
    return total


def correlative_inputs(w, rotational=True, num_points=2048):
    """C3 / C4 inputs at the reference's configured thresholds (pose_graph.lua:20, 40-49). rotational=False
    switches the rotational prefilter off, so that every discrete scan of the sweep is searched (the
    stress variant: ~8 M lowest-resolution candidates for MatchFullSubmap)."""
    synth = w["synth"]
    origin, yaw = w["map_poses"][3]
    rng = np.random.RandomState(4)
    cloud = synth.scan(origin, yaw, 64, 512, w["sph"], rng)
    if num_points > 2048:
        cloud = synth.scan(origin, yaw, 128, 1024, w["sph"], np.random.RandomState(4))
    high = synth.voxel_downsample(cloud, 0.2 if num_points <= 2048 else 0.05, num_points, seed=4)
    low = synth.voxel_downsample(cloud, 0.9, 256, seed=5)
    hist = synth.rotational_histogram(synth.voxel_downsample(cloud, 0.2, 2048, seed=4), HISTOGRAM_SIZE)
    opts = dict(branch_and_bound_depth=7, full_resolution_depth=3, use_rotational_scan_matcher=rotational,
                min_rotational_score=POSE_GRAPH_MIN_ROTATIONAL_SCORE,
                min_low_resolution_score=POSE_GRAPH_MIN_LOW_RESOLUTION_SCORE, linear_xy_search_window=5.0,
                linear_z_search_window=1.0, angular_search_window=float(np.deg2rad(15.0)))
    node = np.concatenate([[0, 0, 0], synth.quat_from_yaw(yaw + 0.4)])
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    return dict(high=high, low=low, hist=hist, submap_hist=submap_histogram(w), opts=opts, node=node,
                submap=submap, origin=origin, yaw=yaw, min_score=POSE_GRAPH_MIN_SCORE)


def local_search_offsets(ci, count, seed=3):
    """C3: node poses inside the local window around the truth (seeded)."""
    from hectorgrapher_b200 import synth
    rng = np.random.RandomState(seed)
    out = []
    for _ in range(count):
        t = ci["origin"] + np.array([rng.uniform(-3, 3), rng.uniform(-3, 3), rng.uniform(-0.6, 0.6)])
        out.append(np.concatenate([t, synth.quat_from_yaw(ci["yaw"] + rng.uniform(-0.2, 0.2))]))
    return out


def cpu_baseline_candidates(w, seconds=8.0, threads=None):
    """CPU restatement of ScoreCandidates on a bounded sample of C4's lowest-resolution lattice."""
    from oracle import pyoracle as o
    threads = threads or os.cpu_count() or 1
    ci = correlative_inputs(w, rotational=False)
    high, hist, node, submap = ci["high"], ci["hist"], ci["node"], ci["submap"]
    hi = o.Grid.tsdf(HI_RES, REL_TRUNC, MAX_WEIGHT)
    hi.set_raw(*w["hi"])
    lo = o.Grid.tsdf(LO_RES, REL_TRUNC, MAX_WEIGHT)
    lo.set_raw(*w["lo"])
    m = o.Fcsm(hi, lo, ci["submap_hist"], **ci["opts"])
    wxy = wz = (hi.grid_size + 1) // 2 + int(np.floor(np.linalg.norm(high, axis=1).max() / HI_RES + 0.5 + 0.5))
    rng = np.random.RandomState(0)
    depth = ci["opts"]["branch_and_bound_depth"] - 1
    def sample(n):
        # lattice positions over the first 8 discrete scans of a narrow yaw window
        return np.stack([rng.randint(0, 8, n), rng.randint(-wxy, wxy + 1, n) // 64 * 64,
                         rng.randint(-wxy, wxy + 1, n) // 64 * 64, rng.randint(-wz, wz + 1, n) // 64 * 64], 1)
    _, _, t = m.score_candidates(wxy, wz, 0.02, node, submap, high, hist, depth, sample(2000), threads=threads,
                                 want_outputs=False)
    n = int(max(2000, min(4000000, 2000 * seconds / max(t, 1e-4))))
    _, _, t = m.score_candidates(wxy, wz, 0.02, node, submap, high, hist, depth, sample(n), threads=threads,
                                 want_outputs=False)
    return n / t, "%d lowest-resolution candidates x %d pts, %d threads" % (n, len(high), threads), threads


def golden():
    try:
        with open(os.path.join(ROOT, "tests", "golden", "c4_golden.json")) as f:
            return json.load(f)
    except Exception:
        return {}


# ncu (profiles/r2_ncu_summary.md) of ScoreLatticePackedKernel<5,1> on the full sweep: the level it gathers from
# is L1-resident, and of the L1 sub-units the LSU data pipe is the busiest (l1tex__data_pipe_lsu_wavefronts 63.5 %
# of its peak of one wavefront = one 128-byte pass per clock and SM; issue slots 69 %). 342.5 M wavefronts for
# 8 022 929 x 2 048 gathers.
LATTICE_L1_WAVEFRONTS_PER_GATHER = 342.5e6 / (8022929 * 2048)
L1_BYTES_PER_CLK_PER_SM = 128
NUM_SMS = 148


def sharded_full_submap(m, ci, rank, world, dist):
    """One MatchFullSubmap over all ranks: phase A on this rank's discrete scans, one all-reduce-max of the
    dive bound (in device memory, on the search's stream), phase B under the common bound, all-gather of the best records and selection — all of it
    inside the caller's timed region. Returns (winner dict or None, order_unresolved, this rank's result)."""
    from hectorgrapher_b200 import dist as hgdist
    handle, d_bound, stream = m.full_submap_begin_async(ci["node"][3:], ci["submap"][3:], ci["high"], ci["low"],
                                                        ci["hist"], min_score=ci["min_score"], shard_index=rank,
                                                        num_shards=world)
    hgdist.all_reduce_max_device_bound(d_bound, stream, dist)  # queued behind phase A: no host wait
    mine = m.full_submap_finish(handle, 0.0)
    best, unresolved = hgdist.reduce_best_candidate(mine, dist)
    return best, unresolved, mine


def correlative_variant(hg, torch, w, rank, world, dist, g_hi, g_lo, rotational, reps=5):
    ci = correlative_inputs(w, rotational=rotational)
    m = hg.Fcsm(g_hi, g_lo, ci["submap_hist"], **ci["opts"])
    for _ in range(2):  # warm-up: sizes the per-thread workspace, first NCCL calls
        sharded_full_submap(m, ci, rank, world, dist)
    torch.cuda.synchronize()
    dts = []
    best = mine = None
    for _ in range(reps):
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        best, unresolved, mine = sharded_full_submap(m, ci, rank, world, dist)
        torch.cuda.synchronize()
        dts.append(time.perf_counter() - t0)
    dt = float(np.median(dts))
    rec = torch.tensor([dt, float(mine["num_scored"]), float(mine["num_scans"]), float(mine["num_lattice"]),
                        float(mine["lattice_kernel_ms"]), float(mine["dive_bound"])], dtype=torch.float64, device="cuda")
    if dist is not None:
        allrec = [torch.zeros_like(rec) for _ in range(world)]
        dist.all_gather(allrec, rec)  # statistics only, after the timed region
        allrec = torch.stack(allrec).cpu().numpy()
    else:
        allrec = rec.cpu().numpy()[None]
    if rank != 0:
        return None
    name = "configured" if rotational else "full_sweep"
    want = golden().get(name)
    checked = None
    if want is not None:
        got = (best["found"], best["score"], best["scan_index"], best["ox"], best["oy"], best["oz"]) if best else (0.0,) * 6
        exp = (want["found"], want["score"], want["scan_index"], want["ox"], want["oy"], want["oz"])
        assert tuple(float(np.float32(v)) for v in got) == tuple(float(np.float32(v)) for v in exp), (
            "C4 %s: selected candidate %r differs from the oracle's %r (tests/golden/c4_golden.json)" % (name, got, exp))
        checked = "equals the CPU oracle's MatchFullSubmap (tests/golden/c4_golden.json)"
    tmax = float(allrec[:, 0].max())
    scored = float(allrec[:, 1].sum())
    lattice_n = float(allrec[:, 3].sum())
    lattice_ms = float(allrec[:, 4].max())
    return {"search_ms": 1e3 * tmax, "candidates_scored": scored, "candidates_per_s": scored / tmax,
            "discrete_scans": float(allrec[:, 2].sum()),
            "best": None if best is None else {"score": float(best["score"]), "scan": int(best["scan_index"]),
                                               "offset": [int(best["ox"]), int(best["oy"]), int(best["oz"])],
                                               "check": checked, "order_unresolved": bool(unresolved)},
            "dive_bound": float(allrec[:, 5].max()),
            "lowest_resolution_kernel": {"candidates": lattice_n, "ms_max_over_ranks": lattice_ms,
                                         "candidates_per_s_per_gpu": (lattice_n / world) / (lattice_ms * 1e-3)
                                         if lattice_ms > 0 else 0.0},
            "num_points": len(ci["high"])}


def correlative_bench(hg, torch, w, rank, world, dist, clocks_mhz=None):
    """M2 on C4: MatchFullSubmap on the 0.10 m TSDF submap at pose_graph.lua's thresholds, discrete scans
    dealt round-robin over the ranks. The timed region is the whole search as a multi-GPU caller runs it:
    per-rank phase A, NCCL all-reduce-max of the dive bound, phase B, NCCL all-gather of the best records and
    the selection. Two variants: 'configured' (rotational prefilter on, as the reference runs) and
    'full_sweep' (prefilter off: every yaw of the sweep is searched, ~8 M lowest-resolution candidates)."""
    g_hi = hg.Grid.tsdf(HI_RES, REL_TRUNC, MAX_WEIGHT, *w["hi"])
    g_lo = hg.Grid.tsdf(LO_RES, REL_TRUNC, MAX_WEIGHT, *w["lo"])
    configured = correlative_variant(hg, torch, w, rank, world, dist, g_hi, g_lo, True)
    sweep = correlative_variant(hg, torch, w, rank, world, dist, g_hi, g_lo, False)
    if rank != 0:
        return None
    n_pts = sweep["num_points"]
    lat = sweep["lowest_resolution_kernel"]
    sm_mhz = clocks_mhz or 1965.0
    l1_peak = NUM_SMS * L1_BYTES_PER_CLK_PER_SM * sm_mhz * 1e6 / 1e9  # GB/s
    l1_achieved = lat["candidates_per_s_per_gpu"] * n_pts * LATTICE_L1_WAVEFRONTS_PER_GATHER * L1_BYTES_PER_CLK_PER_SM / 1e9
    cpu_rate, cpu_desc, cpu_threads = cpu_baseline_candidates(w)
    return {
        "metric": "correlative candidates scored/s", "value": sweep["candidates_per_s"], "unit": "candidates/s",
        "workload": "C4 MatchFullSubmap, 0.10 m TSDF submap, 7-level pyramid (full_resolution_depth 3), N=%d pts, "
                    "min_score %.2f / min_low_resolution_score %.2f (pose_graph.lua:20, 45), discrete scans dealt "
                    "over %d rank(s); timed: phase A + NCCL all-reduce-max of the dive bound + phase B + NCCL "
                    "all-gather of the best records + selection. 'value' is the full sweep (rotational prefilter "
                    "off); 'configured' has it on at min_rotational_score %.2f" % (
                        n_pts, POSE_GRAPH_MIN_SCORE, POSE_GRAPH_MIN_LOW_RESOLUTION_SCORE, world,
                        POSE_GRAPH_MIN_ROTATIONAL_SCORE),
        "search_ms": sweep["search_ms"], "candidates_scored": sweep["candidates_scored"],
        "full_sweep": sweep, "configured": configured,
        "roofline": {"bound": "l1", "kernel": "ScoreLatticePackedKernel<5,true>", "unit": "GB/s",
                     "achieved": l1_achieved, "peak": l1_peak, "frac": l1_achieved / l1_peak,
                     "traffic": measured_traffic("ScoreLatticePackedKernel<5>"),
                     "note": "per GPU, lowest-resolution lattice kernel of the full sweep. The level it gathers "
                             "from is a few hundred KB and L1-resident (ncu: 46 MB of DRAM traffic per 8 M-candidate "
                             "launch, L1 hit rate 96 %%), so HBM is not the bound; the busiest unit is the L1 LSU data "
                             "pipe: achieved = wavefronts (ncu: %.5f per gather; a lane reads four voxels per 32-bit "
                             "load) x 128 B x gathers/s; peak = %d SMs x %d B/clk x %.0f MHz (one wavefront per clock "
                             "and SM). SURVEY 8(d)'s byte model (N + 24 B per candidate against HBM) would read %.2f of "
                             "the HBM peak." % (
                                 LATTICE_L1_WAVEFRONTS_PER_GATHER, NUM_SMS, L1_BYTES_PER_CLK_PER_SM, sm_mhz,
                                 lat["candidates_per_s_per_gpu"] * (n_pts + 24) / 1e9 / measured_peak()[0])},
        "cpu_baseline": {"value": cpu_rate, "unit": "candidates/s", "cores": cpu_threads, "kind": "port",
                         "sample": cpu_desc},
    }


def cpu_baseline_local_search(w, ci, offsets, threads=None):
    """The oracle's FastCorrelativeScanMatcher3D::Match on a few of the same offsets, one search per thread."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import pyoracle as o
    threads = threads or os.cpu_count() or 1
    hi = o.Grid.tsdf(HI_RES, REL_TRUNC, MAX_WEIGHT)
    hi.set_raw(*w["hi"])
    lo = o.Grid.tsdf(LO_RES, REL_TRUNC, MAX_WEIGHT)
    lo.set_raw(*w["lo"])
    m = o.Fcsm(hi, lo, ci["submap_hist"], **ci["opts"])
    offsets = offsets[:max(threads, 4)]
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        rs = list(ex.map(lambda nd: m.match(nd, ci["submap"], ci["high"], ci["low"], ci["hist"],
                                            min_score=ci["min_score"]), offsets))
    dt = time.perf_counter() - t0
    scored = sum(int(r["num_scored"]) for r in rs)
    return {"searches_per_s": len(offsets) / dt, "value": scored / dt, "unit": "candidates/s", "cores": threads,
            "kind": "port", "sample": "%d of the same searches, one per host thread" % len(offsets)}


def local_search_bench(hg, torch, w, rank, world, dist):
    """C3: FastCorrelativeScanMatcher3D::Match (pose_graph.lua:40-49: 7-level pyramid, full_resolution_depth 3,
    5 m / 1 m / 15 deg windows, thresholds 0.55 / 0.77 / 0.55) for 100 seeded node poses inside the window, N = 2048
    points, and 10 at N = 20480; one batched device-resident search per cloud size. Rank 0 only (replicas)."""
    if rank != 0:
        return None
    g_hi = hg.Grid.tsdf(HI_RES, REL_TRUNC, MAX_WEIGHT, *w["hi"])
    g_lo = hg.Grid.tsdf(LO_RES, REL_TRUNC, MAX_WEIGHT, *w["lo"])
    out = {"workload": "C3: local Match, 0.10 m TSDF submap, 7-level pyramid (full_resolution_depth 3), windows 5 m / "
                       "1 m / 15 deg, min_score 0.55, min_rotational_score 0.77, min_low_resolution_score 0.55; seeded "
                       "node poses inside the window, one hg_fcsm_match_batch per cloud size"}
    want = golden().get("c3")
    for n_points, count in ((2048, 100), (20480, 10)):
        ci = correlative_inputs(w, rotational=True, num_points=n_points)
        m = hg.Fcsm(g_hi, g_lo, ci["submap_hist"], **ci["opts"])
        offsets = local_search_offsets(ci, count)
        submaps = [ci["submap"]] * count
        nodes = [(ci["high"], ci["low"], ci["hist"])] * count
        m.match_batch(offsets, submaps, nodes, min_score=ci["min_score"])  # warm-up
        torch.cuda.synchronize()
        dts = []
        for _ in range(3):
            t0 = time.perf_counter()
            rs = m.match_batch(offsets, submaps, nodes, min_score=ci["min_score"])
            dts.append(time.perf_counter() - t0)
        dt = float(np.median(dts))
        scored = sum(r["num_scored"] for r in rs)
        entry = {"points": len(ci["high"]), "searches": count, "found": int(sum(r["found"] for r in rs)),
                 "ms_per_search": 1e3 * dt / count, "searches_per_s": count / dt, "candidates_scored": scored,
                 "candidates_per_s": scored / dt, "discrete_scans_per_search": float(np.mean([r["num_scans"] for r in rs]))}
        if n_points == 2048 and want is not None:
            keys = want["keys"]
            for r, row in zip(rs, want["rows"]):
                got = [float(r[k]) for k in keys]
                assert got[0] == row[0] and (got[0] == 0 or got[:8] == row[:8]), (
                    "C3: search result %r differs from the oracle's %r" % (got, row))
            entry["check"] = "all %d results equal the CPU oracle's (tests/golden/c4_golden.json)" % count
        if n_points == 2048:
            try:
                entry["cpu_baseline"] = cpu_baseline_local_search(w, ci, offsets)
            except Exception as ex:
                entry["cpu_baseline"] = {"error": repr(ex)}
        out["n%d" % n_points] = entry
    return out


C5_FCSM_OPTIONS = dict(branch_and_bound_depth=7, full_resolution_depth=3, use_rotational_scan_matcher=False,
                       min_low_resolution_score=0.3, linear_xy_search_window=5.0, linear_z_search_window=1.0,
                       angular_search_window=float(np.deg2rad(15.0)))
C5_OPTIONS = dict(min_score=0.55, global_localization_min_score=0.6, occupied_space_weights=(5.0, 30.0),
                  translation_weight=10.0, rotation_weight=1.0, max_num_iterations=10)
C5_INSERTER = dict(relative_truncation_distance=REL_TRUNC, maximum_weight=MAX_WEIGHT, num_free_space_voxels=0,
                   epsilon=1.0, sigma=4.0, min_range=0.4, max_range=60.0, insertion_ratio=1.0)


def c5_submap_inputs(synth, sph, s, num_submaps):
    """Submap s of C5: three scans around a pose of the corkscrew (inserted into 0.10 m / 0.45 m TSDF
    grids) and one node scan taken at the pose itself, down-sampled like the adaptive voxel filter does."""
    origin, yaw = synth.corkscrew_pose(s, max(num_submaps, 2))
    scans = []
    for k, d in enumerate(([0.0, 0.0, 0.0], [0.6, -0.4, 0.1], [-0.5, 0.5, -0.1])):
        o = origin + np.array(d)
        c = synth.scan(o, yaw + 0.2 * k, 64, 1024, sph, np.random.RandomState(1000 * s + k))
        q = synth.quat_from_yaw(yaw + 0.2 * k)
        scans.append((o.astype(np.float32), (synth.quat_rotate(q, c.astype(np.float64)) + o).astype(np.float32)))
    cloud = synth.scan(origin, yaw, 64, 512, sph, np.random.RandomState(1000 * s + 7))
    high = synth.voxel_downsample(cloud, 0.2, 2048, seed=6)
    low = synth.voxel_downsample(cloud, 0.9, 256, seed=7)
    return origin, yaw, scans, high, low


def c5_jobs(synth, s, local_index, origin, yaw, high, low, hist, nodes, seed=5):
    rng = np.random.RandomState(seed + 31 * s)
    submap = np.array([0, 0, 0, 1, 0, 0, 0], float)
    jobs = []
    for _ in range(nodes):
        node = np.concatenate([origin + np.array([rng.uniform(-2, 2), rng.uniform(-2, 2), rng.uniform(-0.4, 0.4)]),
                               synth.quat_from_yaw(yaw + rng.uniform(-0.15, 0.15))])
        jobs.append(dict(submap_index=local_index, full=False, node_pose=node, submap_pose=submap, high=high,
                         low=low, hist=hist))
    return jobs


def cpu_baseline_pairs(synth, sph, num_submaps, pairs=16, threads=None):
    """The oracle's FastCorrelativeScanMatcher3D::Match + CeresScanMatcher3D-style refine for 'pairs' jobs of
    submap 0, one pair per host thread (how ConstraintBuilder3D's pool runs them)."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import pyoracle
    threads = threads or os.cpu_count() or 1
    origin, yaw, scans, high, low = c5_submap_inputs(synth, sph, 0, num_submaps)
    g_hi = pyoracle.Grid.tsdf(HI_RES, REL_TRUNC, MAX_WEIGHT)
    g_lo = pyoracle.Grid.tsdf(LO_RES, REL_TRUNC, MAX_WEIGHT)
    for o, pts in scans:
        g_hi.insert_tsdf(o, pts, **C5_INSERTER)
        g_lo.insert_tsdf(o, pts, **C5_INSERTER)
    hist = np.zeros(8, np.float32)
    m = pyoracle.Fcsm(g_hi, g_lo, hist, **C5_FCSM_OPTIONS)
    jobs = c5_jobs(synth, 0, 0, origin, yaw, high, low, hist, pairs)

    def one(j):
        r = m.match(j["node_pose"], j["submap_pose"], high, low, hist, min_score=C5_OPTIONS["min_score"])
        if r["found"]:
            pose = np.array([r["tx"], r["ty"], r["tz"], r["qw"], r["qx"], r["qy"], r["qz"]])
            pyoracle.scan_match([high, low], [g_hi, g_lo], C5_OPTIONS["occupied_space_weights"],
                                C5_OPTIONS["translation_weight"], C5_OPTIONS["rotation_weight"], pose,
                                max_num_iterations=C5_OPTIONS["max_num_iterations"], use_nonmonotonic_steps=False)
        return r["found"], r["num_scored"]

    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        out = list(ex.map(one, jobs))
    dt = time.perf_counter() - t0
    return pairs / dt, threads, "%d pairs of submap 0 (match + refine), %d host threads, %d found" % (
        pairs, threads, sum(int(f) for f, _ in out))


def constraint_bench(hg, torch, w, rank, world, dist, num_submaps=100, nodes_per_submap=100, cpu=True):
    """C5: ConstraintBuilder3D::ComputeConstraint for (node, submap) pairs — local correlative match +
    CeresScanMatcher3D refine. Submaps are partitioned over ranks (the owner holds that submap's grids and
    pyramid); each submap is built on its owner by device-side TSDF insertion."""
    synth, sph = w["synth"], w["sph"]
    hist = np.zeros(8, np.float32)
    empty = np.zeros((0, 3), np.int32)
    mine = [s for s in range(num_submaps) if s % world == rank]
    matchers, g_his, g_los, jobs = [], [], [], []
    for local, s in enumerate(mine):
        origin, yaw, scans, high, low = c5_submap_inputs(synth, sph, s, num_submaps)
        g_hi = hg.Grid.tsdf(HI_RES, REL_TRUNC, MAX_WEIGHT, empty, [], [])
        g_lo = hg.Grid.tsdf(LO_RES, REL_TRUNC, MAX_WEIGHT, empty, [], [])
        for o, pts in scans:
            g_hi.insert_tsdf(o, pts, **C5_INSERTER)
            g_lo.insert_tsdf(o, pts, **C5_INSERTER)
        g_his.append(g_hi)
        g_los.append(g_lo)
        matchers.append(hg.Fcsm(g_hi, g_lo, hist, **C5_FCSM_OPTIONS))
        jobs += c5_jobs(synth, s, local, origin, yaw, high, low, hist, nodes_per_submap)
    batch = hg.ConstraintBatch(matchers, g_his, g_los, jobs, C5_OPTIONS) if mine else None
    if batch is not None:
        batch.run()  # warm-up at full size: every worker thread's scratch memory reaches its steady-state size
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    if batch is not None:
        batch.run()
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    found, scored = batch.totals() if batch is not None else (0, 0)
    stats = torch.tensor([found, scored, len(jobs)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
    if rank != 0:
        return None
    n_jobs = int(stats[2].item())
    out = {"workload": "C5: %d submaps x %d nodes = %d (node, submap) pairs, local Match (5 m / 1 m / 15 deg, "
                       "N=2048) + refine on {high, low} clouds; submaps partitioned over %d rank(s), built by "
                       "device-side TSDF insertion; pairs of a submap searched and refined in groups (one device-resident "
                       "search + one refine launch per group); timed after one untimed pass of the same batch" % (
                           num_submaps, nodes_per_submap, n_jobs, world),
           "pairs_per_s": n_jobs / float(dt.item()), "seconds": float(dt.item()),
           "ms_per_pair_per_gpu": 1e3 * float(dt.item()) * world / max(n_jobs, 1),
           "constraints_found": int(stats[0].item()), "candidates_scored": float(stats[1].item()),
           "candidates_per_s": float(stats[1].item()) / float(dt.item()),
           "group_workers_per_gpu": int(os.environ.get("HG_CONSTRAINT_THREADS", 3))}
    if cpu:
        try:
            rate, threads, desc = cpu_baseline_pairs(synth, sph, num_submaps)
            out["cpu_baseline"] = {"value": rate, "unit": "pairs/s", "cores": threads, "kind": "port", "sample": desc}
        except Exception as ex:
            out["cpu_baseline"] = {"error": repr(ex)}
    return out


def c1_bench(hg, torch, w):
    """C1 (BASELINE.json configs[0], the reference's own CPU-runnable case): one 32 x 640 scan against a
    5 cm TSDF submap (+ 0.45 m low-resolution grid for the two-cloud Match). Evaluate through the
    ceres::CostFunction surface with host pointers, and a full CeresScanMatcher3D::Match; the oracle
    on one thread beside it (the reference runs each scan matcher with num_threads = 1)."""
    from oracle import pyoracle
    synth, sph = w["synth"], w["sph"]
    empty = np.zeros((0, 3), np.int32)
    g_hi = hg.Grid.tsdf(0.05, REL_TRUNC, MAX_WEIGHT, empty, [], [])
    g_lo = hg.Grid.tsdf(LO_RES, REL_TRUNC, MAX_WEIGHT, empty, [], [])
    o_hi = pyoracle.Grid.tsdf(0.05, REL_TRUNC, MAX_WEIGHT)
    o_lo = pyoracle.Grid.tsdf(LO_RES, REL_TRUNC, MAX_WEIGHT)
    origin, yaw = w["map_poses"][2]
    for k, d in enumerate(([0.0, 0.0, 0.0], [0.4, -0.3, 0.1], [-0.3, 0.4, -0.1])):
        o = origin + np.array(d)
        c = synth.scan(o, yaw + 0.1 * k, 32, 640, sph, np.random.RandomState(300 + k))
        pts = (synth.quat_rotate(synth.quat_from_yaw(yaw + 0.1 * k), c.astype(np.float64)) + o).astype(np.float32)
        for g in (g_hi, g_lo, o_hi, o_lo):
            g.insert_tsdf(o.astype(np.float32), pts, **C5_INSERTER)
    rng = np.random.RandomState(1)
    cloud = synth.scan(origin, yaw, 32, 640, sph, rng)
    low = synth.voxel_downsample(cloud, 0.9, 1024, seed=3)
    start = np.concatenate([origin + rng.uniform(-0.1, 0.1, 3), synth.quat_from_yaw(yaw + rng.uniform(-0.05, 0.05))])
    scale = 1.0 / np.sqrt(len(cloud))
    cost = hg.Cost("tsdf", scale, cloud, [g_hi])
    for _ in range(5):
        cost.evaluate(start)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 100
    for _ in range(reps):
        cost.evaluate(start)
    eval_ms = 1e3 * (time.perf_counter() - t0) / reps
    for _ in range(3):
        hg.scan_match([cloud, low], [g_hi, g_lo], [5.0, 30.0], 10.0, 1.0, start, max_num_iterations=12,
                      use_nonmonotonic_steps=False)
    t0 = time.perf_counter()
    for _ in range(20):
        pose, summary = hg.scan_match([cloud, low], [g_hi, g_lo], [5.0, 30.0], 10.0, 1.0, start,
                                      max_num_iterations=12, use_nonmonotonic_steps=False)
    match_ms = 1e3 * (time.perf_counter() - t0) / 20
    ocost = pyoracle.Cost("tsdf", scale, cloud, [o_hi])
    ocost.evaluate(start)
    t0 = time.perf_counter()
    for _ in range(3):
        ocost.evaluate(start)
    cpu_eval_ms = 1e3 * (time.perf_counter() - t0) / 3
    t0 = time.perf_counter()
    opose = pyoracle.scan_match([cloud, low], [o_hi, o_lo], [5.0, 30.0], 10.0, 1.0, start, max_num_iterations=12,
                                use_nonmonotonic_steps=False)[0]
    cpu_match_ms = 1e3 * (time.perf_counter() - t0)
    return {"workload": "C1: one 32 x 640 scan (%d points) against a 5 cm TSDF submap; Evaluate = "
                        "ceres::CostFunction::Evaluate with host pointers (residuals + 2 Jacobian blocks back over "
                        "PCIe), Match = CeresScanMatcher3D::Match on {scan, 1024-point low-resolution cloud}" % len(cloud),
            "evaluate_ms": eval_ms, "evaluate_points_per_s": len(cloud) / (eval_ms * 1e-3),
            "match_ms": match_ms, "match_iterations": int(summary[2]),
            "pose_difference_to_cpu": float(np.abs(pose[:3] - opose[:3]).max()),
            "cpu_baseline": {"evaluate_ms": cpu_eval_ms, "match_ms": cpu_match_ms, "cores": 1, "kind": "port",
                             "sample": "oracle Jet functor / oracle cost functions under the same LM, 1 thread"}}


def insertion_bench(hg, torch, w, scans=6):
    """N1: TSDFRangeDataInserter3D::Insert of whole scans (100k returns each) into a device-resident
    0.10 m TSDF submap, host points in, grid ready for matching out; CPU baseline = the oracle's
    sequential inserter (the reference inserts on one thread). Rank 0 only: one submap does not shard."""
    from oracle import pyoracle
    synth = w["synth"]
    clouds = []
    for k in range(scans):
        origin, yaw = w["map_poses"][k % len(w["map_poses"])]
        rng = np.random.RandomState(100 + k)
        c = synth.scan(origin, yaw, 64, 1600, w["sph"], rng)
        q = synth.quat_from_yaw(yaw)
        clouds.append((origin.astype(np.float32),
                       (synth.quat_rotate(q, c.astype(np.float64)) + origin).astype(np.float32)))
    opts = dict(relative_truncation_distance=2.5, maximum_weight=1000.0, num_free_space_voxels=0, epsilon=1.0,
                sigma=4.0, min_range=0.4, max_range=60.0, insertion_ratio=1.0)
    empty = np.zeros((0, 3), np.int32)
    g = hg.Grid.tsdf(HI_RES, REL_TRUNC, MAX_WEIGHT, empty, [], [])
    for origin, pts in clouds[:2]:  # warm-up: allocations, first bricks
        g.insert_tsdf(origin, pts, **opts)
    g = hg.Grid.tsdf(HI_RES, REL_TRUNC, MAX_WEIGHT, empty, [], [])
    g.insert_tsdf(*clouds[0], **opts)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    updates = 0
    returns = 0
    for origin, pts in clouds[1:]:
        updates += g.insert_tsdf(origin, pts, **opts)
        returns += len(pts)
    torch.cuda.synchronize()
    gpu_s = time.perf_counter() - t0
    o = pyoracle.Grid.tsdf(HI_RES, REL_TRUNC, MAX_WEIGHT)
    o.insert_tsdf(*clouds[0], **opts)
    t0 = time.perf_counter()
    for origin, pts in clouds[1:3]:
        o.insert_tsdf(origin, pts, **opts)
    cpu_s = time.perf_counter() - t0
    cpu_returns = sum(len(p) for _, p in clouds[1:3])
    return {"workload": "N1: %d scans x ~%d returns inserted into a 0.10 m TSDF submap (default ray cast, "
                        "truncation 2.5 voxels), host points in, apron + decoded copies regenerated after every "
                        "scan" % (scans - 1, returns // max(1, scans - 1)),
            "returns_per_s": returns / gpu_s, "cell_updates_per_s": updates / gpu_s,
            "ms_per_scan": 1e3 * gpu_s / (scans - 1), "bricks": int(g.num_bricks),
            "cpu_baseline": {"value": cpu_returns / cpu_s, "unit": "returns/s", "cores": 1, "kind": "port",
                             "sample": "2 of the same scans, sequential oracle inserter"}}


def bind_to_gpu_numa_node(torch, local_rank):
    """One process per GPU: run on (and allocate pinned host memory from) the NUMA node the GPU hangs off,
    so that the end-to-end copies of several ranks do not all cross the socket interconnect. Best effort:
    a box without NUMA information (or a single node) is left alone. Returns a description for the log."""
    try:
        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = torch.cuda.get_device_properties(local_rank).pci_domain_id
        dev = torch.cuda.get_device_properties(local_rank).pci_device_id
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bus, dev)
        node = int(open(path).read().strip())
        if node < 0:
            return "no NUMA node reported for the GPU"
        cpus = []
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus += list(range(int(lo), int(hi or lo) + 1))
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if not allowed:
            return "NUMA node %d has no CPU this process may use" % node
        os.sched_setaffinity(0, allowed)
        return "bound to NUMA node %d (%d CPUs)" % (node, len(allowed))
    except Exception as ex:  # no sysfs, old torch, containers without the files ...
        return "not bound (%s)" % type(ex).__name__


def run_ours(args, rank, world, local_rank):
    import torch
    import hectorgrapher_b200 as hg
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    hg.lib()
    torch.cuda.set_device(local_rank)
    hg.set_device(local_rank)
    if world > 1:
        log("[rank %d] %s" % (rank, bind_to_gpu_numa_node(torch, local_rank)))
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist = dist_mod
    t_setup = time.time()
    w = build_workload()
    N = int(w["offsets"][-1])
    g_hi = hg.Grid.tsdf(HI_RES, REL_TRUNC, MAX_WEIGHT, *w["hi"])
    g_lo = hg.Grid.tsdf(LO_RES, REL_TRUNC, MAX_WEIGHT, *w["lo"])
    batch = hg.Batch("interpolated_multires_tsdf", w["points"], w["offsets"], w["scales"], [g_hi, g_lo], w["ratios"])
    log("[rank %d] setup %.1f s: %d pts, grids %d + %d voxels, %d + %d bricks" % (
        rank, time.time() - t_setup, N, len(w["hi"][0]), len(w["lo"][0]), g_hi.num_bricks, g_lo.num_bricks))

    dev = torch.device("cuda", local_rank)
    d_r = torch.empty(N, dtype=torch.float64, device=dev)
    d_jt0 = torch.empty(N * 3, dtype=torch.float64, device=dev)
    d_jq0 = torch.empty(N * 4, dtype=torch.float64, device=dev)
    d_jt1 = torch.empty(N * 3, dtype=torch.float64, device=dev)
    d_jq1 = torch.empty(N * 4, dtype=torch.float64, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    poses = np.ascontiguousarray(w["poses"])
    i0, i1 = w["i0"], w["i1"]
    stream = torch.cuda.current_stream()

    def step_device():
        batch.evaluate_raw(poses.ctypes.data, len(poses), i0.ctypes.data, i1.ctypes.data, True, d_r.data_ptr(),
                           d_jt0.data_ptr(), d_jq0.data_ptr(), d_jt1.data_ptr(), d_jq1.data_ptr(),
                           stream.cuda_stream)

    for _ in range(max(args.warmup, 3)):
        step_device()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    launches0 = hg.kernel_launch_count()
    times = []
    torch.cuda.synchronize()
    for _ in range(args.steps):
        flush.fill_(1)  # evict L2 between timed iterations (256 MiB > 126 MB L2); not timed
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        step_device()
        e1.record(stream)
        e1.synchronize()
        times.append(e0.elapsed_time(e1))
    torch.cuda.synchronize()
    launches = hg.kernel_launch_count() - launches0
    if dist is not None:
        dist.barrier()
    total_ms = torch.tensor([float(np.sum(times))], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())

    # sanity: finite outputs, non-trivial Jacobian
    chk = float(d_r.abs().sum().item())
    jchk = float(d_jq1.abs().max().item())
    assert np.isfinite(chk) and jchk > 0, "bench outputs are not plausible"

    # end to end through the C-ABI with HOST buffers: this step's scans H2D from pinned memory,
    # Evaluate, residuals + Jacobians D2H into pinned host arrays.
    h_pts = torch.from_numpy(w["points"]).pin_memory()
    h_out = [torch.empty(n, dtype=torch.float64).pin_memory() for n in (N, 3 * N, 4 * N, 3 * N, 4 * N)]
    def step_e2e():
        batch.update_points(h_pts.numpy())
        batch.evaluate_raw(poses.ctypes.data, len(poses), i0.ctypes.data, i1.ctypes.data, False,
                           h_out[0].data_ptr(), h_out[1].data_ptr(), h_out[2].data_ptr(), h_out[3].data_ptr(),
                           h_out[4].data_ptr(), None)
    step_e2e()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    e2e_steps = max(2, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    torch.cuda.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e_s.item())
    clocks = sampler.stop() if sampler else None  # sampled across the device-timed and end-to-end loops
    assert abs(float(h_out[0].abs().sum()) - chk) <= 1e-9 * max(1.0, chk)

    # Gauss-Newton surface (no Jacobian materialised), for context
    batch.normal_equations(poses, i0, i1)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        batch.normal_equations(poses, i0, i1)
    ne_ms = 1e3 * (time.perf_counter() - t0) / 3

    corr = None
    try:
        corr = correlative_bench(hg, torch, w, rank, world, dist,
                                 clocks_mhz=(clocks or {}).get("sm_mhz") if clocks else None)
    except Exception as ex:  # the headline must not be lost to the secondary metric
        log("[rank %d] correlative bench failed: %r" % (rank, ex))
        corr = {"error": repr(ex)}
    if dist is not None:
        # every rank must leave the correlative section together, whatever happened in it
        dist.barrier()
    local = None
    try:
        local = local_search_bench(hg, torch, w, rank, world, dist)
    except Exception as ex:
        log("[rank %d] local search bench failed: %r" % (rank, ex))
        local = {"error": repr(ex)}
    cons = None
    try:
        if args.c5_small:
            cons = constraint_bench(hg, torch, w, rank, world, dist, num_submaps=8, nodes_per_submap=24)
        else:
            cons = constraint_bench(hg, torch, w, rank, world, dist)
    except Exception as ex:
        log("[rank %d] constraint bench failed: %r" % (rank, ex))
        cons = {"error": repr(ex)}

    ins = None
    c1 = None
    if rank == 0:
        try:
            ins = insertion_bench(hg, torch, w)
        except Exception as ex:
            log("[rank 0] insertion bench failed: %r" % (ex,))
            ins = {"error": repr(ex)}
        try:
            c1 = c1_bench(hg, torch, w)
        except Exception as ex:
            log("[rank 0] C1 bench failed: %r" % (ex,))
            c1 = {"error": repr(ex)}

    if rank == 0:
        peak, peak_src = measured_peak()
        ms_per_step = total_ms / args.steps
        value = world * N / (ms_per_step * 1e-3)
        per_gpu = N / (ms_per_step * 1e-3)
        achieved = per_gpu * ALG_BYTES_PER_POINT / 1e9
        cpu_rate, cpu_desc, cpu_threads, _ = cpu_baseline_points(w)
        line = {
            "metric": "scan-to-TSDF residual+Jacobian points/s", "value": value, "unit": "points/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C2 multi-resolution TSDF registration: 64 scans x 100k pts, 0.10 m + 0.45 m "
                                   "TSDF, two-pose interpolated multi-res functor, 8 control poses; one step = one "
                                   "batched Evaluate (residuals + 4 Jacobian blocks, Ceres layout)",
                       "points_per_step_per_gpu": N, "multi_gpu": "replicas only (one window per GPU)",
                       "l2": "256 MiB flush write between timed steps; working set per step ~0.9 GB"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": measured_traffic("CostKernelPersistent<2,true>"),
                         "traffic_note": "bytes per launch (ncu, profiles/r2_ncu_summary.md); algorithmic bytes per "
                                         "launch = %d" % (ALG_BYTES_PER_POINT * N),
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_point": ALG_BYTES_PER_POINT, "kernel": "CostKernelPersistent<2,true>"},
            "cpu_baseline": {"value": cpu_rate, "unit": "points/s", "cores": cpu_threads, "kind": "port",
                             "sample": cpu_desc},
            "e2e": {"value": world * N * e2e_steps / e2e_s, "unit": "points/s",
                    "h2d_bytes_per_step": int(w["points"].nbytes + poses.nbytes),
                    "d2h_bytes_per_step": int(15 * N * 8), "ms_per_step": 1e3 * e2e_s / e2e_steps},
            "gpu_launches": int(launches), "clocks": clocks,
            "normal_equations": {"ms_per_step": ne_ms, "points_per_s": N / (ne_ms * 1e-3),
                                 "note": "hg_batch_normal_equations: host poses in, 64 x (14x14 JtJ, Jtr, cost) out"},
            "correlative": corr,
            "local_search": local,
            "constraint_batch": cons,
            "insertion": ins,
            "c1_scan_to_submap": c1,
        }
        emit(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    # Libraries (NCCL's version banner, for one) may print to stdout; the contract is ONE JSON line
    # there. Route fd 1 to stderr for the whole run and keep the real stdout for that line.
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--c5-small", action="store_true", help="C5 on 8 x 24 pairs instead of 100 x 100 (quick runs)")
    ap.add_argument("--c5-full", action="store_true",
                    help="run the constraint batch at BASELINE.json's full C5 size (100 submaps x 100 nodes)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
