#!/usr/bin/env python
"""bench.py - the EKF-SLAM hot path on B200: scan-updates/s at N = 10,000 landmarks.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--landmarks N] [--m M]

One "step" = one scan update of ONE filter: pose injection + prediction + m x (Mahalanobis association, gain,
state update) + the fused covariance pass (BASELINE.json metric; configs[3] "single large map N=10,000
landmarks (n=20,003), 20 measurements/scan").  Prints ONE JSON line (contract in the task statement):

  value         scans/s with inputs resident in HBM (ekf_scan_dev), CUDA events on the kernels' stream, no
                per-scan events inside the timed region
  e2e           the same through the host-buffer C-ABI call ekf_scan (H2D + D2H inside the timed region)
  roofline      ekf_downdate_kernel (HBM-bound): algorithmic bytes / its mean launch time, taken live from a
                second pass of K scans with per-scan events (ekf_enable_timing); `others` = the N = 60,000 pass
                and the batched kernel
  cpu_baseline  oracle B ("port": oracle/ekf_struct.c) on all host cores, WHOLE scans (every one of the m
                measurements) continued from the GPU filter's own state, with the GPU/CPU parity of those very
                scans (association identical, state, Sigma norm-wise and on the observed block) and the decision
                margins; `others` = the batched configuration on a sample of filters, and the reference's own
                compiled slam.cpp ("reference": oracle/_ref/libslam_ref.so) at its native N = 50
  run           how this run was parallelised / what was checked (kept out of `config`, which is identical in
                both arms)

N > 1 GPUs: the SAME single map, Sigma row-sharded over the ranks (equal-area split, replicated sequential
phase; per scan the owners store the speculated column panel straight into every rank's buffer over NVLink peer
memory - `shard_push_cols_kernel` + arrival flags, no collective on the per-scan path):
strong scaling, also at N = 1.  Sub-objects: `n60k` = BASELINE configs[4] (N = 60,000; one GPU holds the 57.7 GB
triangle, N GPUs row-shard it), `batched` = configs[2] (65,536 independent 20-landmark filters PER GPU, one warp
each, no communication: weak scaling).

--impl reference: the reference's CPU algorithm on the host cores (oracle B; every measurement of every step), same
config, metric and unit; rank 0 only.
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
for _p in (ROOT, os.path.join(ROOT, "oracle")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "ekf_scan_updates_per_s_N10k"
UNIT = "scan-updates/s"
NOMINAL_HBM_GBS = 8000.0          # the "about 8 TB/s" of BASELINE.json's north star
FP64_PEAK_TFLOPS = 37.1           # DMMA m8n8k4, measured on this pool: profiles/r1_fp64_peak_microbench.txt


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--landmarks", type=int, default=10000)
    ap.add_argument("--m", type=int, default=20)
    ap.add_argument("--seed", type=int, default=7)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-batched", action="store_true", help="skip the batched-filters section")
    ap.add_argument("--no-n60k", action="store_true", help="skip the N = 60,000 section (BASELINE configs[4])")
    ap.add_argument("--no-small", action="store_true", help="skip the N = 50 section (reference's native size)")
    ap.add_argument("--overlap", action="store_true",
                    help="single GPU: run the covariance pass of a scan behind the next scan's sequential side "
                         "(ekf_set_overlap; measured SLOWER at N=10k, DESIGN.md section 10 - kept for reproduction)")
    ap.add_argument("--batch-filters", type=int, default=65536, help="filters per GPU (BASELINE configs[2])")
    ap.add_argument("--batch-landmarks", type=int, default=20)
    ap.add_argument("--batch-m", type=int, default=20)
    ap.add_argument("--batch-steps", type=int, default=10)
    ap.add_argument("--n60k-steps", type=int, default=10)
    ap.add_argument("--cpu-scans", type=int, default=6, help="whole scans the cpu_baseline leg runs at N=10k")
    ap.add_argument("--cpu-batch-filters", type=int, default=2048, help="filters of the batched cpu_baseline sample")
    return ap.parse_args()


def peaks():
    """measured roofline denominators written by the driver, else the recipe's fallback"""
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def config_of(N, m):
    """the workload, identically worded in both arms"""
    return {"workload": f"single large map N={N} landmarks (n={3 + 2 * N}), {m} measurements/scan, unknown "
                        "association, all landmarks mapped (BASELINE.json configs[3])",
            "landmarks": N, "m": m, "map": "ekf_fill_procedural(seed) / oracle/procedural_map.py: the same map in both arms",
            "l2": f"inputs larger than L2: the covariance triangle ({8.0 * (3 + 2 * N) * (4 + 2 * N) / 2 / 1e9:.2f} GB) "
                  "is streamed once per scan"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index: int):
        self.dev = device_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.dev), "-lms", "25"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons, power = [], None, set(), []
        for (t, line) in self.rows:
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                if t0 <= t <= t1 + 0.12:
                    sm.append(float(parts[1])); power.append(float(parts[3]))
                smax = float(parts[2])
            except ValueError:
                continue
            if t0 <= t <= t1 + 0.12:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                                   parts[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax,
                "power_w": float(np.median(power)) if power else None, "samples": len(sm),
                "reasons": sorted(reasons)}


def _sm_clock(device_index):
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm", "--format=csv,noheader,nounits", "-i",
                              str(device_index)], capture_output=True, text=True, timeout=10).stdout.strip()
        return float(out.splitlines()[0])
    except Exception:
        return None


def make_workload(state0, count, m, n_scans, seed):
    """Per-scan poses and sightings for the large-map configs: the robot circles in the middle of the landmark
    grid; each scan sights the m nearest landmarks (body frame, sigma = 1e-3 per axis like nusim's fake sensor,
    nusim.cpp:88,202,339-340)."""
    rng = np.random.default_rng(seed)
    lm = state0[3:3 + 2 * count].reshape(-1, 2)
    side = np.sqrt(count)
    cx = cy = side / 2.0
    rad = max(1.5, side / 5.0)
    poses = np.zeros((n_scans, 3))
    xy = np.zeros((n_scans, m, 2))
    ids = np.zeros((n_scans, m), dtype=np.int32)
    for s in range(n_scans):
        ang = 0.004 * s
        x, y = cx + rad * np.cos(ang), cy + rad * np.sin(ang)
        th = ang + np.pi / 2
        d = np.hypot(lm[:, 0] - x, lm[:, 1] - y)
        near = np.argpartition(d, m)[:m]
        rng.shuffle(near)
        c, sn = np.cos(th), np.sin(th)
        dx, dy = lm[near, 0] - x, lm[near, 1] - y
        xy[s] = np.stack([c * dx + sn * dy, -sn * dx + c * dy], axis=1) + rng.normal(0, 1e-3, (m, 2))
        ids[s] = near
        poses[s] = (th, x, y)
    return poses, xy, ids


def algorithmic_bytes(n, m):
    """B_scan = 8 n (n+1) + 8 n 9 m; the covariance pass alone = 8 n (n+1)."""
    return 8.0 * n * (n + 1) + 8.0 * n * 9 * m, 8.0 * n * (n + 1)


def oracle_threads():
    import struct_oracle as so
    L = so.lib()
    L.ekfo_num_threads.restype = int
    return int(L.ekfo_num_threads())


def cpu_filter_from_generator(N, seed):
    """oracle B holding the SAME procedural map the GPU arm fills on the device (oracle/procedural_map.py is the
    host re-implementation of ekf_fill_procedural)"""
    import struct_oracle as so
    from procedural_map import ProceduralMap
    side = int(np.ceil(np.sqrt(N)))
    pm = ProceduralMap(seed, N, N, side)
    o = so.StructEkf(n_landmarks=N)
    o.state[:] = pm.state()
    for i0 in range(0, o.n, 2048):
        i1 = min(o.n, i0 + 2048)
        o.sigma[i0:i1, :] = pm.sigma0_rows(i0, i1)
    o.count = N
    return o


# ----------------------------------------------------------------------------------------------- reference arm
def run_reference(args):
    """--impl reference: the reference's CPU algorithm on the host cores, same config, metric and unit.  slam.cpp
    itself is compiled here (oracle/_ref/libslam_ref.so) but is O(n^3) per measurement and hard-wired dense: hours
    per scan at N = 10,000, so the arm times oracle B, the structured restatement that the tests
    pin to it; the `native_n50` entry times the reference's own object code at its native size."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    N, m = args.landmarks, args.m
    o = cpu_filter_from_generator(N, args.seed)
    cores = oracle_threads()
    total = args.warmup + args.steps
    poses, xy, ids = make_workload(np.array(o.state), N, m, total, args.seed + 100)
    o.set_prev(poses[0][1], poses[0][2])
    secs, ok = [], True
    for i in range(total):
        t0 = time.perf_counter()
        o.predict(*poses[i])
        a = o.update_unknown(xy[i])
        secs.append(time.perf_counter() - t0)
        ok = ok and bool((np.asarray(a) == ids[i]).all())
    o.close()
    secs = secs[args.warmup:]
    value = float(len(secs) / sum(secs))
    desc = (f"oracle B (oracle/ekf_struct.c: structured O(n^2)-per-measurement restatement of slam.cpp:321-489, pthreads "
            f"on all host cores) at N={N}: every step is one whole scan (prediction + all {m} measurements); "
            f"{args.steps} step(s) after {args.warmup} warm-up")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 / value, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_of(N, m),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc,
                             "association_equals_expected_ids": ok},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if not args.no_small:
        line["cpu_baseline"]["others"] = [native_n50_reference()]
    _emit(_OUT, json.dumps(line))


def small_trace():
    """BASELINE configs[1] stand-in at the reference's native size (N = 50, n = 103): circle drive, detections
    only, shuffled, Mahalanobis association"""
    from ekf_slam_b200 import synth
    return [s for s in synth.circle_drive(n_scans=120, seed=1, drop_out_of_range=True, shuffle=True) if len(s.xy)]


def native_n50_reference():
    """the reference's own object code (slam.cpp compiled unmodified into oracle/_ref/libslam_ref.so) on its
    native configuration, one host core"""
    import slam_ref
    if not slam_ref.available():
        return {"config": "N=50", "unavailable": "oracle/_ref/libslam_ref.so not built (needs /root/reference)"}
    scans = small_trace()
    f = slam_ref.RefSlam(n_landmarks=50)
    t0 = time.perf_counter()
    for s in scans:
        f.set_odometry(*s.odom)
        f.sensor_cb([(x, y) for (x, y) in s.xy])
    dt = time.perf_counter() - t0
    return {"config": "N=50 (n=103), unknown association, circle drive around 4 landmarks (BASELINE configs[0]/[1] shape)",
            "value": len(scans) / dt, "unit": UNIT, "cores": 1, "kind": "reference",
            "sample": f"Slam::sensor_cb of the reference's slam.cpp compiled unmodified against the stub Armadillo "
                      f"(oracle/stubs/armadillo: plain triple-loop products, no BLAS), {len(scans)} scans"}


# ----------------------------------------------------------------------------------------------- batched filters
def bench_batched(args, torch, ekf, stream, local, world, dist):
    """BASELINE config 3: independent small filters, one warp each, sharded over the ranks with no communication
    (weak scaling: `--batch-filters` filters PER GPU)."""
    from ekf_slam_b200 import synth
    B, N, m = args.batch_filters, args.batch_landmarks, args.batch_m
    rank = int(os.environ.get("RANK", "0"))
    K, W = args.batch_steps, 4
    g = ekf.EkfBatch(B, n_landmarks=N, device=local)
    g.set_stream(stream.cuda_stream)
    fields = synth.batch_fields(B, N, seed=args.seed + 17 * rank)
    rng = np.random.default_rng(args.seed + 1000 + rank)
    T = W + 2 * K
    h_pose, h_xy, d_pose, d_xy, ids = [], [], [], [], []
    for s in range(T):
        p = synth.batch_pose(B, s)
        xy, order = synth.batch_scan(fields, p, rng, m=N if s < 2 else m)   # the first scans map every landmark
        h_pose.append(p); h_xy.append(xy); ids.append(order)
        d_pose.append(torch.from_numpy(p).cuda()); d_xy.append(torch.from_numpy(xy).cuda())
    d_assoc = torch.zeros((T, B, N), dtype=torch.int32, device="cuda")
    d_cnt = torch.zeros(B, dtype=torch.int32, device="cuda")

    def step(s):
        g.step_dev(d_pose[s].data_ptr(), d_xy[s].shape[1], d_xy[s].data_ptr(), 0, 0, d_assoc[s].data_ptr(),
                   d_cnt.data_ptr(), 0)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for s in range(W):
        step(s)
    l0 = g.launch_count()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for s in range(W, W + K):
        step(s)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = g.launch_count() - l0
    # the landmark numbering of a filter is the order of its first scan: ids[0]; expected association of
    # a later sighting of true landmark L is the position of L in that order
    pos = np.empty_like(ids[0]); np.put_along_axis(pos, ids[0], np.arange(N, dtype=np.int32)[None].repeat(B, 0), axis=1)
    ok = True
    for s in (W, W + K - 1):
        exp = np.take_along_axis(pos, ids[s], axis=1)
        ok = ok and bool((d_assoc[s, :, :ids[s].shape[1]].cpu().numpy() == exp).all())
    ok = ok and g.last_status() == 0 and int(d_cnt.min()) == N

    # end to end: inputs in pinned host memory (as a producer would leave them), results into pinned host arrays
    def pin(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        return t, t.numpy()
    keep = []
    p_in = {}
    for s in range(W + K, W + 2 * K):
        tp, ap = pin(h_pose[s]); tx, ax = pin(h_xy[s])
        keep += [tp, tx]; p_in[s] = (ap, ax)
    t_a = torch.empty((B, m), dtype=torch.int32).pin_memory()
    t_c = torch.empty((B,), dtype=torch.int32).pin_memory()
    t_p = torch.empty((B, 3), dtype=torch.float64).pin_memory()
    out = (t_a.numpy(), t_c.numpy(), t_p.numpy())
    # (the filters are in steady state - all landmarks mapped - so a sighting set may be replayed as warm-up)
    for i in range(3):
        s = W + K + (i % K)
        g.step(p_in[s][0], p_in[s][1], out=out)
    barrier()
    t0 = time.perf_counter()
    for s in range(W + K, W + 2 * K):
        g.step(p_in[s][0], p_in[s][1], out=out)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1000.0
    ok = ok and int(out[1].min()) == N
    tt = torch.tensor([ms, e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(tt[0]), float(tt[1])
    byt = g.step_bytes(m)
    peak, _ = peaks()
    value = world * B * K / (ms / 1000.0)
    res = {"metric": "ekf_batched_filter_steps_per_s", "value": value, "unit": "filter-steps/s",
           "filters_per_gpu": B, "landmarks": N, "m": m, "steps": K, "ms_per_step": ms / K, "scaling": "weak",
           "association_bit_exact_vs_expected_ids": ok, "gpu_launches": int(launches),
           "bytes_per_filter_step": byt, "hbm_frac": (value / world * byt / 1e9) / peak,
           "e2e": {"value": world * B * K / (e2e_ms / 1000.0), "unit": "filter-steps/s",
                   "h2d_bytes_per_step": B * (24 + 16 * m), "d2h_bytes_per_step": B * (4 * m + 4 + 24) + 4}}
    if rank == 0 and not args.no_cpu_baseline:
        res["cpu_baseline"] = batched_cpu_baseline(args, g, fields, h_pose, h_xy, ids, W, K)
    g.close()
    return res


def batched_cpu_baseline(args, g, fields, h_pose, h_xy, ids, W, K):
    """oracle B over a sample of the SAME filters (continued from the GPU filters' own state after the timed
    steps), all host cores: one Python thread per core, each stepping its share of the sample (the C calls
    release the GIL; a 43-dim filter never spawns the oracle's own worker threads)."""
    import struct_oracle as so
    from concurrent.futures import ThreadPoolExecutor
    cores = os.cpu_count() or 1
    Bs = min(args.cpu_batch_filters, args.batch_filters)
    N, m = args.batch_landmarks, args.batch_m
    S = 3                                     # steps per sampled filter
    steps = list(range(W + K, W + K + S))
    filt = []
    for i in range(Bs):
        o = so.StructEkf(n_landmarks=N)
        o.state[:] = g.state(i); o.sigma[:] = g.sigma(i); o.count = g.count(i)
        # filter_previous_configuration = the pose the filter ended its last scan with
        o.set_prev(float(o.state[1]), float(o.state[2]))
        filt.append(o)
    ok = [True] * Bs

    def work(lo, hi):
        for i in range(lo, hi):
            o = filt[i]
            for s in steps:
                o.predict(*h_pose[s][i])
                a = o.update_unknown(h_xy[s][i])
                ok[i] = ok[i] and len(a) == m
    chunks = [(Bs * c // cores, Bs * (c + 1) // cores) for c in range(cores)]
    t0 = time.perf_counter()
    with ThreadPoolExecutor(cores) as ex:
        list(ex.map(lambda c: work(*c), chunks))
    dt = time.perf_counter() - t0
    for o in filt:
        o.close()
    return {"value": Bs * S / dt, "unit": "filter-steps/s", "cores": cores, "kind": "port",
            "sample": f"oracle B on {Bs} of the {args.batch_filters} filters x {S} steps (N={N}, m={m}), continued from "
                      f"the GPU filters' state, {cores} threads", "all_scans_completed": all(ok)}


# ----------------------------------------------------------------------------------------------- one large map
class MapRun:
    """one large map on this process' GPU (or this rank's row shard of it)"""

    def __init__(self, args, torch, dist, ekf, N, m, stream, local, world, rank):
        self.torch, self.dist, self.world, self.rank = torch, dist, world, rank
        self.N, self.m, self.n = N, m, 3 + 2 * N
        self.sharded = world > 1
        side = int(np.ceil(np.sqrt(N)))
        if self.sharded:
            from ekf_slam_b200 import shard
            self.drv = shard.ShardedEkf(N, rank, world, local, stream=stream, max_scan=m)
            self.f = self.drv.h
            self.f.fill_procedural(args.seed, N, side)       # same seed everywhere: every rank fills its own rows
            self.drv.sync()
        else:
            self.drv = None
            self.f = ekf.EkfSlam(n_landmarks=N, device=local, max_scan=m)
            self.f.set_stream(stream.cuda_stream)
            self.f.fill_procedural(args.seed, N, side)
        self.stream = stream
        self.stored, self.n_pad = self.f.storage()

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def close(self):
        (self.drv if self.sharded else self.f).close()

    def run(self, args, K, W, local, overlap=False, clocks=True):
        torch, dist, f, drv, m, N = self.torch, self.dist, self.f, self.drv, self.m, self.N
        overlap = (not self.sharded) and overlap and f.set_overlap(True)
        state0 = f.state()
        n_scans = W + 3 * K                    # warm-up | timed | per-kernel timing | end to end
        poses, xy, ids = make_workload(state0, N, m, n_scans, args.seed + 100)
        f.set_previous_xy(poses[0][1], poses[0][2])
        d_pose = torch.from_numpy(poses).cuda()
        d_xy = torch.from_numpy(xy).cuda()
        d_assoc = torch.full((n_scans, m), -9, dtype=torch.int32, device="cuda")
        d_done = torch.full((n_scans,), m, dtype=torch.int32, device="cuda")

        def scan_dev(s):
            if self.sharded:
                drv.scan_dev(d_pose[s], d_xy[s], d_assoc[s], check=False, done_t=d_done[s:s + 1])
            else:
                f.scan_dev(d_pose[s].data_ptr(), m, d_xy[s].data_ptr(), 0, 0, d_assoc[s].data_ptr())

        # ---- device-resident throughput (`value`): no per-scan events in here -----------------------------
        for s in range(W):
            scan_dev(s)
        l0 = f.launch_count()
        sampler = ClockSampler(local)
        if clocks:
            sampler.start()
            time.sleep(0.3)
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(self.stream)
        for s in range(W, W + K):
            scan_dev(s)
        if overlap:
            f.set_overlap(False)      # drain: `stream` waits for the last pass, so e1 closes over ALL the work
        e1.record(self.stream)
        self.barrier()
        t1 = time.perf_counter()
        clk = sampler.stop(t0, t1) if clocks else None
        ms_total = e0.elapsed_time(e1)
        launches = f.launch_count() - l0
        status = f.last_status()
        if overlap:
            f.set_overlap(True)
        # ---- per-kernel durations: the same K-scan loop again with the handle's per-scan events ------------
        f.enable_timing(K)
        for s in range(W + K, W + 2 * K):
            scan_dev(s)
        if overlap:
            f.set_overlap(False)
        self.barrier()
        seq_ms, dd_ms = f.read_timing(K)
        xa_ms, xp_ms = f.read_timing_shard(K) if self.sharded else (np.zeros(0), np.zeros(0))
        f.enable_timing(0)
        if overlap:
            f.set_overlap(True)
        # ---- end to end through the host-buffer call (`e2e`) ----------------------------------------------
        e_assoc = []
        self.barrier()
        t2 = time.perf_counter()
        for s in range(W + 2 * K, W + 3 * K):
            if self.sharded:
                a, cnt, _, _ = drv.scan(poses[s], xy[s])
            else:
                a, cnt, _ = f.scan(poses[s], xy[s])
            e_assoc.append(a)
        self.barrier()
        e2e_s = time.perf_counter() - t2

        tt = torch.tensor([ms_total, e2e_s * 1000.0], dtype=torch.float64, device="cuda")
        if self.world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total, e2e_ms = float(tt[0]), float(tt[1])
        assoc_dev = d_assoc[W:W + 2 * K].cpu().numpy()
        assoc_ok = bool((assoc_dev == ids[W:W + 2 * K]).all() and bool((d_done[:W + 2 * K] == m).all()) and
                        all((np.asarray(a) == ids[W + 2 * K + i]).all() for i, a in enumerate(e_assoc)) and status == 0)
        ok = torch.tensor([1 if assoc_ok else 0], device="cuda")
        stored_all = torch.tensor([float(self.stored)], dtype=torch.float64, device="cuda")
        if self.world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            dist.all_reduce(stored_all, op=dist.ReduceOp.SUM)
        # per-rank view of the same loop (the sequential kernel of a rank also waits for the slowest rank's columns)
        per_rank = None
        if self.world > 1:
            mine = {"seq_ms": float(np.mean(seq_ms)) if len(seq_ms) else None,
                    "pass_ms": float(np.mean(dd_ms)) if len(dd_ms) else None,
                    "assoc_ms": float(np.mean(xa_ms)) if len(xa_ms) else None,
                    "push_ms": float(np.mean(xp_ms)) if len(xp_ms) else None,
                    "sm_mhz": (clk or {}).get("sm_mhz") if self.rank == 0 else _sm_clock(local)}
            gathered = [None] * self.world
            dist.all_gather_object(gathered, mine)
            per_rank = {k: [g[k] for g in gathered] for k in mine}
        b_scan, b_pass = algorithmic_bytes(self.n, m)
        peak, peak_src = peaks()
        dd_avg_ms = float(np.mean(dd_ms)) if len(dd_ms) else float("nan")
        # the covariance pass of this rank streams its share of the triangle: algorithmic bytes per launch
        b_pass_rank = b_pass * self.stored / float(stored_all.item())
        achieved = b_pass_rank / (dd_avg_ms * 1e-3) / 1e9
        flops_rank = 2.0 * m * self.n * (self.n + 1) * self.stored / float(stored_all.item())
        value = K / (ms_total / 1000.0)
        self.poses, self.xy, self.ids, self.next_scan = poses, xy, ids, W + 3 * K
        return {"value": value, "ms_per_step": ms_total / K, "e2e_value": K / (e2e_ms / 1000.0),
                "e2e_ms_per_step": e2e_ms / K, "launches": int(launches), "assoc_ok": bool(ok.item()), "clocks": clk,
                "overlap": bool(overlap), "per_rank": per_rank,
                "roofline": {"bound": "hbm", "kernel": "ekf_downdate_kernel", "achieved": achieved, "peak": peak,
                             "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                             "frac_nominal_8tbs": achieved / NOMINAL_HBM_GBS,
                             "algorithmic_bytes_per_launch": b_pass_rank, "avg_launch_ms": dd_avg_ms,
                             "launch_time_source": f"handle CUDA events around every launch of a second pass of {K} scans "
                                                   "(the timed region itself carries no per-scan events)",
                             "fp64_tflops_achieved": flops_rank / (dd_avg_ms * 1e-3) / 1e12,
                             "fp64_tflops_peak": FP64_PEAK_TFLOPS,
                             "fp64_peak_source": "DMMA m8n8k4 microbenchmark on this pool (profiles/r1_fp64_peak_microbench.txt)",
                             "seq_phase_avg_ms": float(np.mean(seq_ms)) if len(seq_ms) else None,
                             "pass_share_of_step": dd_avg_ms / (ms_total / K),
                             "scan_bytes": b_scan,
                             "scan_frac_of_hbm_roofline": (b_scan * value / 1e9) / (peak * self.world),
                             "scan_frac_of_nominal_8tbs": (b_scan * value / 1e9) / (NOMINAL_HBM_GBS * self.world)}}


def parity_and_cpu_baseline(args, run):
    """The cpu_baseline leg at N = 10,000: oracle B continues from the GPU filter's own state and runs WHOLE scans;
    the GPU runs the same scans.  Returns the baseline object, with the parity of those scans and the decision
    margins oracle B saw."""
    import struct_oracle as so
    f, N, m, n = run.f, run.N, run.m, run.n
    S = max(1, args.cpu_scans)
    o = so.StructEkf(n_landmarks=N)
    o.state[:] = f.state()
    for i0 in range(0, n, 2048):
        i1 = min(n, i0 + 2048)
        o.sigma[i0:i1, :] = f.sigma_block(i0, i1, 0, n)
    o.count = N
    o.set_prev(*f.previous_xy())
    s0 = run.next_scan - S                    # re-uses the tail of the workload (the filter is in steady state)
    secs, same, best, second, touched = [], True, [], [], set()
    thr = 2.0
    for s in range(s0, s0 + S):
        t0 = time.perf_counter()
        o.predict(*run.poses[s])
        a, b1, b2 = o.update_unknown_margins(run.xy[s])
        secs.append(time.perf_counter() - t0)
        ag, _, _ = f.scan(run.poses[s], run.xy[s])
        same = same and list(ag) == list(a)
        best += list(b1); second += list(b2)
        touched.update(int(k) for k in a)
    best, second = np.array(best), np.array(second)
    # parity of the two filters after those scans: state, Sigma norm-wise over sampled full-width block rows, and
    # the observed block (pose + every landmark sighted in these scans)
    st = f.state()
    state_rel = float(np.linalg.norm(st - o.state) / np.linalg.norm(o.state))
    worst = 0.0
    for i0 in list(np.linspace(0, n - 256, 12).astype(int)):
        dev = f.sigma_block(int(i0), int(i0) + 256, 0, n)
        ref = o.sigma[int(i0):int(i0) + 256, :]
        worst = max(worst, float(np.linalg.norm(dev - ref) / np.linalg.norm(ref)))
    idx = [0, 1, 2] + [3 + 2 * k + e for k in sorted(touched) for e in (0, 1)]
    dev = np.stack([f.sigma_block(i, i + 1, 0, n)[0, idx] for i in idx])
    ref = np.array(o.sigma)[np.ix_(idx, idx)] if n <= 4096 else np.stack([np.asarray(o.sigma[i, :])[idx] for i in idx])
    active_rel = float(np.linalg.norm(dev - ref) / np.linalg.norm(ref))
    o.close()
    value = S / sum(secs)
    return {"value": value, "unit": UNIT, "cores": oracle_threads(), "kind": "port",
            "sample": f"oracle B (oracle/ekf_struct.c, pthreads on all host cores) at N={N}: {S} WHOLE scans (prediction + "
                      f"all {m} measurements each), continued from the GPU filter's own state and Sigma",
            "parity_of_these_scans": {"association_identical": bool(same), "state_rel": state_rel,
                                      "sigma_rel_normwise_worst_block_row": worst,
                                      "sigma_rel_observed_block": active_rel,
                                      "observed_block_dim": len(idx)},
            "decision_margins": {"source": "oracle B, every measurement of these scans",
                                 "min_rel_gap_best_to_threshold": float(np.min(np.abs(best - thr) / thr)),
                                 "min_rel_gap_best_to_second": float(np.min((second - best) / np.maximum(best, 1e-300))),
                                 "measurements": int(len(best))}}


def small_gpu_vs_reference(ekf, local):
    """N = 50 (the reference's native size): the same trace through ekf_scan and through the reference's compiled
    slam.cpp - throughput of both and the parity of the final filters"""
    import slam_ref
    from ekf_slam_b200 import synth
    ref = native_n50_reference()
    if "unavailable" in ref:
        return ref
    scans = small_trace()
    r = slam_ref.RefSlam(n_landmarks=50)
    g = ekf.EkfSlam(n_landmarks=50, device=local)
    host = synth.MapHost()
    same = True
    t_gpu = 0.0
    for s in scans:
        r.set_odometry(*s.odom)
        ar = r.sensor_cb([(x, y) for (x, y) in s.xy])
        x, y, th = host.composed(s.odom)
        t0 = time.perf_counter()
        ag, _, pose = g.scan((th, x, y), s.xy)
        t_gpu += time.perf_counter() - t0
        host.after_update(tuple(pose), s.odom)
        same = same and list(ag) == list(ar)
    sg, sr = g.state(), r.state
    SG, SR = g.sigma(), r.sigma
    k = 3 + 2 * r.counter_obstacles
    ref["gpu"] = {"value": len(scans) / t_gpu, "unit": UNIT, "path": "ekf_scan, host buffers, one scan per call (latency-bound)"}
    ref["parity_vs_reference_object_code"] = {
        "association_identical": bool(same),
        "state_rel": float(np.linalg.norm(sg - sr) / np.linalg.norm(sr)),
        "sigma_rel_normwise": float(np.linalg.norm(SG - SR) / np.linalg.norm(SR)),
        "sigma_rel_observed_block": float(np.linalg.norm(SG[:k, :k] - SR[:k, :k]) / np.linalg.norm(SR[:k, :k]))}
    g.close()
    return ref


def run_ours(args):
    import torch
    import torch.distributed as dist
    import ekf_slam_b200 as ekf

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the product has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    N, m, K, W = args.landmarks, args.m, args.steps, args.warmup
    # a non-default stream: the legacy default stream's handle is NULL, which ekf_set_stream reads as
    # "use the handle's own stream" - the timing events must sit on the stream the kernels use
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0

    run = MapRun(args, torch, dist, ekf, N, m, stream, local, world, rank)
    res = run.run(args, K, W, local, overlap=args.overlap)
    roof = res["roofline"]
    prof = os.path.join(ROOT, "profiles", "downdate_traffic.json")
    if os.path.exists(prof) and world == 1 and N == 10000 and m == 20:
        with open(prof) as fh:
            tj = json.load(fh)
        roof["traffic"] = tj.get("dram_bytes_per_launch")
        roof["traffic_source"] = ("committed ncu --set full capture of this kernel at this size ("
                                  + str(tj.get("source", "profiles/")) + "), not measured in this run")
    cpu = None
    if not args.no_cpu_baseline and world == 1 and rank == 0:
        cpu = parity_and_cpu_baseline(args, run)
    stored_gb = 8.0 * run.n * (run.n + 1) / 2 / 1e9
    n_pad = run.n_pad
    run.close()
    del run
    torch.cuda.empty_cache()

    n60k = None
    if not args.no_n60k:
        free, _ = torch.cuda.mem_get_info()
        need = 8.0 * 120003 * 120004 / 2 / world * 1.08 + 2e9
        if free < need:
            n60k = {"unavailable": f"needs {need / 1e9:.0f} GB of free HBM per GPU, {free / 1e9:.0f} GB free"}
        else:
            K6 = args.n60k_steps
            r6 = MapRun(args, torch, dist, ekf, 60000, m, stream, local, world, rank)
            x6 = r6.run(args, K6, 2, local, clocks=False)
            r6.close()
            del r6
            torch.cuda.empty_cache()
            n60k = {"metric": "ekf_scan_updates_per_s_N60k", "value": x6["value"], "unit": UNIT, "n_gpus": world,
                    "steps": K6, "warmup": 2, "ms_per_step": x6["ms_per_step"], "scaling": "strong",
                    "config": "single huge map N=60,000 landmarks (n=120,003; 57.7 GB stored triangle), 20 "
                              "measurements/scan (BASELINE.json configs[4])" + (f", row-sharded over {world} GPUs" if world > 1 else ""),
                    "association_bit_exact_vs_expected_ids": x6["assoc_ok"],
                    "e2e": {"value": x6["e2e_value"], "unit": UNIT, "ms_per_step": x6["e2e_ms_per_step"]},
                    "gpu_launches": x6["launches"], "roofline": x6["roofline"]}

    batched = None
    if not args.no_batched:
        batched = bench_batched(args, torch, ekf, stream, local, world, dist)
    small = None
    if not args.no_small and rank == 0 and not args.no_cpu_baseline:
        small = small_gpu_vs_reference(ekf, local)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    others = []
    if n60k is not None and "roofline" in n60k:
        r6 = n60k["roofline"]
        others.append({"kernel": "ekf_downdate_kernel", "config": "N=60000 (BASELINE configs[4])", "bound": "hbm",
                       "achieved": r6["achieved"], "peak": r6["peak"], "unit": "GB/s", "frac": r6["frac"],
                       "avg_launch_ms": r6["avg_launch_ms"], "seq_phase_avg_ms": r6["seq_phase_avg_ms"],
                       "scan_updates_per_s": n60k["value"], "e2e_scan_updates_per_s": n60k["e2e"]["value"]})
    if batched is not None:
        others.append({"kernel": "ekf_batch_kernel", "config": f"{batched['filters_per_gpu']} filters/GPU x N={batched['landmarks']}, "
                                                               f"m={batched['m']} (BASELINE configs[2])",
                       "bound": "fp64 issue / latency (not HBM: see DESIGN.md section 6)", "unit": "GB/s",
                       "achieved": batched["hbm_frac"] * roof["peak"], "peak": roof["peak"], "frac": batched["hbm_frac"],
                       "filter_steps_per_s": batched["value"], "e2e_filter_steps_per_s": batched["e2e"]["value"]})
    roof["others"] = others
    line = {
        "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_of(N, m),
        "run": {"parallelism": (f"one map, Sigma row-sharded over {world} GPUs (equal-area split); per scan the owners "
                                f"store the {m}x{n_pad}x2 speculated column panel into every rank's buffer over NVLink peer "
                                "memory (shard_push_cols_kernel + arrival flags), no collective on the per-scan path")
                               if world > 1 else "1 GPU",
                "overlap": ("covariance pass of scan s runs out of place on a second stream behind the sequential "
                            "side of scan s+1 (ekf_set_overlap); the timed region ends after the last pass")
                           if res["overlap"] else "off (pass and sequential phase back to back on one stream)",
                "association_bit_exact_vs_expected_ids": res["assoc_ok"],
                "stored_triangle_gb": stored_gb, "per_rank": res["per_rank"]},
        "clocks": res["clocks"],
        "e2e": {"value": res["e2e_value"], "unit": UNIT, "h2d_bytes_per_step": 24 + 16 * m, "d2h_bytes_per_step": 32 + 4 * m,
                "ms_per_step": res["e2e_ms_per_step"]},
        "gpu_launches": res["launches"],
        "roofline": roof,
    }
    if n60k is not None:
        line["n60k"] = n60k
    if batched is not None:
        line["batched"] = batched
    if cpu is not None:
        oth = []
        if batched is not None and "cpu_baseline" in batched:
            oth.append(dict(batched["cpu_baseline"], config="batched (BASELINE configs[2])"))
        if small is not None:
            oth.append(small)
        cpu["others"] = oth
        line["cpu_baseline"] = cpu
    if world > 1:
        dist.destroy_process_group()
    _emit(_OUT, json.dumps(line))


def _quiet_stdout():
    """The contract is ONE JSON line on stdout.  Libraries (NCCL prints its version banner to stdout) must
    not add to it: everything written to fd 1 while the benchmark runs goes to stderr instead; the JSON
    line is written to the saved descriptor."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    return saved


def _emit(saved_fd, text):
    sys.stdout.flush()
    os.write(saved_fd, (text + "\n").encode())


if __name__ == "__main__":
    a = parse_args()
    _OUT = _quiet_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
