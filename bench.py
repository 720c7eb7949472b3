#!/usr/bin/env python
"""bench.py -- throughput of the per-step adversarial input path (BASELINE.json metric):
perturb (scale, clamp, add) -> hard | dynamic voxelize -> HardSimpleVFE | scatter mean forward -> backward to
the raw per-point perturbation, in M input points/s, on synthetic sweeps of the named shapes.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload kitti|nuscenes|nuscenes-dynamic|kitti-dynamic] [--batch B] [--points N]

A *step* is one pass of the path over one batch (kitti: B=6 sweeps x 120k points x 4 features =
BASELINE.json configs[1], the default; nuscenes: B=4 x 300k x 5 = configs[2]; nuscenes-dynamic = configs[3];
--points/--batch = the configs[4] sweep).  N > 1 is launched by torchrun, one rank per GPU, every rank its
own batch (weak scaling, no data-path collective), and the step ends with the NCCL all-reduce of the
perturber's flat parameter-gradient buffer (34 641 fp32 for KITTI).

Timing: every step is bracketed by its own pair of CUDA events on the launching stream; between steps,
outside the event pairs, a 512 MiB buffer is overwritten so that every step starts with a cold L2; the K
step times are summed, MAX over ranks.  `value` = inputs resident in HBM, CUDA-graph replay.  `e2e` = the
public host front end (StepPipeline): pinned host buffers in and out, every step's H2D and D2H copies inside
the timed region, three streams, double-buffered.  `roofline` = per-kernel CUDA-event durations measured
inside the library (rpc_profile_*), algorithmic bytes per launch, measured HBM peak; `traffic` = DRAM bytes per
launch from the committed ncu capture of the same workload (profiles/).  `--impl reference` times the CPU
oracle (mmcv CPU semantics, dense per-call table -- see oracle/rpc_oracle.c) on the host cores instead.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PERTURBER_PARAMS = {4: 34641, 5: 34772}        # hidden [64,128,64]
L2_FLUSH_BYTES = 512 << 20
POINTS_OVERRIDE = 0                            # --points


def ncu_traffic(args, kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture of this workload at its default
    size (profiles/traffic.json, written by tools/ncu_traffic.py); null for any other size."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if args.batch or args.points or not os.path.exists(path):
        return {}
    with open(path) as f:
        e = json.load(f).get("%s:%s" % (args.workload, kernel))
    return {"traffic": e["bytes"], "traffic_source": e["source"]} if e else {}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def b_alg_hard(N, C, P, M, K, F):
    """SURVEY 8(d) algorithmic bytes of the whole hard step (sum over the batch)"""
    return 4 * (N * C + M * P * C + 5 * M + 3 * K * C + 2 * M * F)


class ClockSampler(threading.Thread):
    """polls SM clock and throttle reasons through NVML while the timed region runs"""

    def __init__(self, index: int, period_s: float = 0.005):
        super().__init__(daemon=True)
        self.index, self.period = index, period_s
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            pass

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {getattr(nv, k): k for k in dir(nv) if k.startswith("nvmlClocksThrottleReason")
                 or k.startswith("nvmlClocksEventReason")}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if isinstance(bit, int) and bit and (r & bit) and bin(bit).count("1") == 1:
                        self.reasons.add(nm.replace("nvmlClocksThrottleReason", "").replace("nvmlClocksEventReason", ""))
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        benign = {"None", "GpuIdle", "ApplicationsClocksSetting", "All"}
        bad_map = {"HwSlowdown": "hw_slowdown", "HwThermalSlowdown": "hw_thermal_slowdown",
                   "SwThermalSlowdown": "sw_thermal_slowdown", "SwPowerCap": "sw_power_cap",
                   "HwPowerBrakeSlowdown": "hw_power_brake", "SyncBoost": "sync_boost",
                   "DisplayClockSetting": "display_clock"}
        reasons = sorted({bad_map.get(r, r) for r in self.reasons if r not in benign})
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(s)}


def workload(name: str, batch: int = 0):
    """(preset, sweeps per GPU, dynamic?)  kitti = configs[1], nuscenes = configs[2], nuscenes-dynamic = configs[3]"""
    from robustpointclouds_b200 import synth
    dynamic = name.endswith("-dynamic")
    base = name.split("-")[0]
    p = dict(synth.PRESETS[base])
    if POINTS_OVERRIDE:
        p["N"] = POINTS_OVERRIDE
    B = batch or (6 if base == "kitti" else 4)
    return p, B, dynamic


def b_alg_dyn(N, C, M, F):
    """SURVEY 8(d) algorithmic bytes of the whole dynamic step"""
    return 4 * (3 * N * C + 4 * N + 2 * M * F + 4 * M)


def cpu_step(dynamic, cloud, raw, eb, fb, p, gf, buf):
    """one sweep of the path on the CPU oracle (the timed unit of the CPU legs)"""
    from oracle import ref_path as orc
    C, P, V = p["C"], p["max_num_points"], p["max_voxels"][0]
    if not dynamic:
        return orc.step_fused(cloud, raw, eb, fb, p["voxel_size"], p["point_cloud_range"], P, V, C, gf,
                              table_mode=0, buf=buf)[0]
    y, _ = orc.perturb_apply(cloud, raw, eb, fb)
    co = orc.dynamic_voxelize(y, p["voxel_size"], p["point_cloud_range"])
    f, oc, mp, cnt = orc.dynamic_scatter_mean(y, co)
    g = orc.dynamic_scatter_mean_bwd(gf[: len(cnt)], mp, cnt)
    orc.perturb_bwd(g, raw, eb, fb)
    return len(cnt)


# ------------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's CPU implementation of the path (oracle restatement of mmcv's CPU kernels, dense
    per-call cell table like mmcv) on the host cores, one sweep per thread."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from concurrent.futures import ThreadPoolExecutor

    from oracle import ref_path as orc
    from robustpointclouds_b200 import synth

    p, B, dynamic = workload(args.workload, args.batch)
    base = args.workload.split("-")[0]
    C, P, V = p["C"], p["max_num_points"], p["max_voxels"][0]
    vs, pcr = p["voxel_size"], p["point_cloud_range"]
    cores = len(os.sched_getaffinity(0))
    threads = max(1, min(cores, 32))            # every thread allocates its own dense table (360 MB for KITTI)
    clouds = synth.make_batch(base, B, N=p["N"])
    raws = [synth.make_raw(c, 99 + i) for i, c in enumerate(clouds)]
    eb, fb = orc.perturber_bounds(C, True)
    gf = np.random.default_rng(0).normal(size=(max(V, p["N"]) if dynamic else V, C)).astype(np.float32)
    orc.lib()
    # a step = S sweeps drawn round-robin from the batch, one per host thread, S bounded so that the
    # whole run ends within ~2 minutes
    bufs = [orc.StepBuffers(len(clouds[0]), C, P, V, C)]
    t0 = time.perf_counter()
    cpu_step(dynamic, clouds[0], raws[0], eb, fb, p, gf, bufs[0])
    t1 = time.perf_counter() - t0
    S = threads
    if (args.steps + args.warmup) * t1 * 1.5 > 120.0:      # even one sweep per thread per step is too long
        args.steps = max(3, int(120.0 / (t1 * 1.5)) - args.warmup)
    bufs = [orc.StepBuffers(len(clouds[i % B]), C, P, V, C) for i in range(S)]

    def one(i):
        j = i % B
        return cpu_step(dynamic, clouds[j], raws[j], eb, fb, p, gf, bufs[i])

    idx = list(range(S))
    with ThreadPoolExecutor(threads) as ex:
        for _ in range(args.warmup):
            list(ex.map(one, idx))
        t0 = time.perf_counter()
        for _ in range(args.steps):
            list(ex.map(one, idx))
        dt = time.perf_counter() - t0
    pts = sum(len(clouds[i % B]) for i in idx) * args.steps
    val = pts / dt / 1e6
    sample = (f"{S} sweeps per step drawn round-robin from the batch's {B}, {threads} threads (one sweep per thread), " +
              ("dynamic voxelize + sort-unique scatter-mean + backward (oracle port)" if dynamic else
               f"dense int32[{'x'.join(map(str, orc.grid_size(vs, pcr)[::-1]))}] table allocated+filled per call like mmcv"))
    line = {
        "impl": "reference", "metric": "M points/s perturb+voxelize+VFE fwd/bwd", "value": round(val, 4),
        "unit": "M points/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(dt / args.steps * 1e3, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args.workload, p, B, note="CPU oracle port of the reference's mmcv CPU path"),
        "cpu_baseline": {"value": round(val, 4), "unit": "M points/s", "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": round(val, 4), "unit": "M points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def config_dict(name, p, B, note=None, **kw):
    dynamic = name.endswith("-dynamic")
    wl = {"kitti": "SECOND KITTI adversarial step: perturb(0.2 m)+hard voxelize+HardSimpleVFE fwd/bwd",
          "nuscenes": "CenterPoint nuScenes 10-sweep hard step: perturb+hard voxelize+HardSimpleVFE fwd/bwd",
          "nuscenes-dynamic": "CenterPoint nuScenes 10-sweep dynamic step (no caps): perturb+dynamic voxelize+"
                              "scatter-mean VFE fwd/bwd",
          "kitti-dynamic": "KITTI dynamic step (no caps): perturb+dynamic voxelize+scatter-mean VFE fwd/bwd"}[name]
    d = {"workload": f"{wl}, batch {B} x {p['N']} pts x {p['C']}",
         "voxel_size": p["voxel_size"], "point_cloud_range": p["point_cloud_range"],
         "max_num_points": -1 if dynamic else p["max_num_points"],
         "max_voxels": -1 if dynamic else p["max_voxels"][0], "batch_per_gpu": B,
         "points_per_sweep": p["N"], "distribution": "lidar (synth.py)"}
    if note:
        d["note"] = note
    d.update(kw)
    return d


def cpu_baseline(name, clouds, raws, p, budget_s=12.0):
    """oracle (port) on ONE host core, bounded sample of the same batch"""
    from oracle import ref_path as orc
    dynamic = name.endswith("-dynamic")
    C, P, V = p["C"], p["max_num_points"], p["max_voxels"][0]
    eb, fb = orc.perturber_bounds(C, True)
    gf = np.random.default_rng(0).normal(size=(max(V, max(len(c) for c in clouds)) if dynamic else V, C)).astype(np.float32)
    buf = orc.StepBuffers(max(len(c) for c in clouds), C, P, V, C)
    orc.lib()
    done, pts = 0, 0
    t0 = time.perf_counter()
    while True:
        i = done % len(clouds)
        b = orc.StepBuffers(len(clouds[i]), C, P, V, C) if len(clouds[i]) != buf.perturbed.shape[0] else buf
        cpu_step(dynamic, clouds[i], raws[i], eb, fb, p, gf, b)
        done += 1
        pts += len(clouds[i])
        dt = time.perf_counter() - t0
        if dt > budget_s or done >= 4 * len(clouds):
            break
    return {"value": round(pts / dt / 1e6, 4), "unit": "M points/s", "cores": 1, "kind": "port",
            "sample": f"{done} sweeps of the batch, sequential on 1 core, {dt:.1f} s, " +
                      ("sort-unique scatter-mean" if dynamic else "dense per-call cell table like mmcv")}


# ------------------------------------------------------------------------------------------------
def pgrad_pattern(n, rank, dev):
    """flat perturber-gradient stand-in of rank r: r + 1 + i * 2^-10 -- exact in fp32, and so is every partial sum
    over <= 8 ranks, so the all-reduced buffer can be compared with the closed form bit for bit"""
    import torch
    return (torch.arange(n, dtype=torch.float32, device=dev) * (1.0 / 1024.0) + float(rank + 1)).contiguous()


def pgrad_expected(n, world, dev):
    import torch
    return torch.arange(n, dtype=torch.float32, device=dev) * (world / 1024.0) + float(world * (world + 1) // 2)


class Collective:
    """the step's only exchange: all-reduce of the perturber's flat parameter gradients (the reference gets
    it from DDP, custom_hook.py:24-28), through this library's NVLink peer-memory kernel when every rank can map its
    peers, else NCCL -- agreed by all ranks"""

    def __init__(self, args, world, rank, dev):
        import torch
        import torch.distributed as dist
        self.world, self.rank, self.dev, self.kind = world, rank, dev, args.allreduce
        self.sync_par = None
        self.sync_buf = torch.zeros(4, dtype=torch.float32, device=dev)
        self._pars = {}
        self.peer = False
        if world > 1 and args.allreduce in ("ll", "p2p"):
            from robustpointclouds_b200 import PeerAllReduce
            ok = 1.0
            try:
                self.sync_par = PeerAllReduce(4, dev)
            except Exception as e:                      # no peer mapping on this box: every rank falls back together
                print(f"[bench] rank {rank}: peer all-reduce unavailable ({type(e).__name__}: {e}); using NCCL", file=sys.stderr)
                ok = 0.0
            agree = torch.tensor([ok], device=dev)
            dist.all_reduce(agree, op=dist.ReduceOp.MIN)
            self.peer = agree.item() >= 1.0
            if not self.peer:
                self.sync_par = None

    def for_size(self, n):
        """PeerAllReduce for a buffer of n floats (None: NCCL / single GPU)"""
        if not self.peer:
            return None
        if n not in self._pars:
            from robustpointclouds_b200 import PeerAllReduce
            self._pars[n] = PeerAllReduce(n, self.dev, mode="ll" if self.kind == "ll" else "oneshot")
        return self._pars[n]

    def rank_sync(self):
        """device-side rendezvous of the ranks, OUTSIDE the per-step event pairs: every rank flushes its L2 at its
        own pace, and without this the skew is paid inside the timed step by the all-reduce waiting for the peer"""
        import torch.distributed as dist
        if self.sync_par is not None:
            self.sync_par.all_reduce(self.sync_buf)
        elif self.world > 1:
            dist.all_reduce(self.sync_buf)

    def describe(self, n):
        if self.world == 1:
            return "none"
        how = ("nccl all_reduce" if not self.peer else
               "low-latency NVLink peer-memory all-reduce (rpc_allreduce_ll: value + call number pushed as one 8-byte store, in the step's graph)"
               if self.kind == "ll" else "one-shot NVLink peer-memory all-reduce (rpc_allreduce_oneshot, in the step's graph)")
        return how + " of %d fp32 perturber grads per step" % n

    def timeouts(self):
        t = 0
        for q in list(self._pars.values()) + ([self.sync_par] if self.sync_par is not None else []):
            t += int(q.state[2].item())
        return t


def measure(args, name, B_override, points, steps, warmup, coll, rank, world, dev, full):
    """One workload on this rank's GPU: device-resident step time (CUDA-graph replay, cold L2), end-to-end time through
    StepPipeline (pinned host buffers), per-kernel durations and the roofline figures.  `full`: the headline workload
    (all legs); otherwise a shorter e2e leg and no lean / cpu legs."""
    import torch
    import torch.distributed as dist

    from robustpointclouds_b200 import PerturbCfg, StepPipeline, VoxelCfg, _lib, ops, synth
    from robustpointclouds_b200.voxel_perturber import perturber_bounds

    global POINTS_OVERRIDE
    saved = POINTS_OVERRIDE
    POINTS_OVERRIDE = points
    p, B, dynamic = workload(name, B_override)
    POINTS_OVERRIDE = saved
    C, P, V = p["C"], p["max_num_points"], p["max_voxels"][0]
    F = C
    cfg = VoxelCfg.make(p["voxel_size"], p["point_cloud_range"], P, V, C, F)
    clouds = synth.make_batch(name.split("-")[0], B, N=p["N"], seed=1234 + rank * B)
    raws = [synth.make_raw(c, 99 + rank * B + i) for i, c in enumerate(clouds)]
    n = [len(c) for c in clouds]
    N = sum(n)
    eb, fb = perturber_bounds(C, True)                      # the product's own table (no oracle in the GPU arm)
    bounds = PerturbCfg.make(eb.tolist(), fb.tolist())
    h_pts = torch.from_numpy(np.concatenate(clouds)).pin_memory()
    h_raw = torch.from_numpy(np.concatenate(raws)).pin_memory()
    pts, raw = h_pts.to(dev), h_raw.to(dev)
    # the graph-captured step owns its workspace (rpc_voxelize_hard keeps it clean between calls)
    ws = torch.empty(((ops.dynamic_workspace_bytes if dynamic else ops.hard_workspace_bytes)(cfg, n) + 256,),
                     dtype=torch.uint8, device=dev)
    if dynamic:
        out = ops.dynamic_scatter_mean(pts, n, cfg, raw=raw, bounds=bounds, workspace=ws)
        cap = N
    else:
        out = ops.hard_voxelize(pts, n, cfg, raw=raw, bounds=bounds, want_feats=True, want_maps=True, want_slots=False,
                                workspace=ws)
        cap = out.capacity
    gfe = torch.randn((cap, F), device=dev, generator=torch.Generator(dev).manual_seed(5 + rank))
    graw = torch.empty((N, C), dtype=torch.float32, device=dev)
    npar = PERTURBER_PARAMS.get(C, 34641)
    pgrad = pgrad_pattern(npar, rank, dev)                  # this rank's flat perturber gradients (a fixed pattern)
    psum = torch.zeros_like(pgrad)                          # their sum over the ranks
    par = coll.for_size(npar)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)
    flush_rd = torch.zeros(L2_FLUSH_BYTES // 8, dtype=torch.int32, device=dev) if args.flush == "write+read" else None
    flush_sink = torch.zeros(1, dtype=torch.int64, device=dev)

    def l2_flush():
        """overwrite 512 MiB (evicts everything), then -- with --flush write+read -- stream 256 MiB of clean data
        through L2 so that the step does not also pay for writing back the flush's own dirty lines"""
        flush.zero_()
        if flush_rd is not None:
            flush_sink.copy_(flush_rd.sum())

    def fwd():
        if dynamic:
            ops.dynamic_scatter_mean(pts, n, cfg, raw=raw, bounds=bounds, out=out, workspace=ws)
        else:
            ops.hard_voxelize(pts, n, cfg, raw=raw, bounds=bounds, out=out, workspace=ws)

    def bwd():
        ops.step_bwd(gfe, out.num_points, out.point2voxel, cfg, raw=raw, bounds=bounds, out=graw)

    def path():
        fwd()
        bwd()

    def exchange():                                     # perturber-gradient all-reduce: psum = sum over ranks of pgrad
        if par is not None:
            par.all_reduce(pgrad, psum)
        elif world > 1:
            psum.copy_(pgrad)
            dist.all_reduce(psum)

    graph = None
    if not args.no_graph:
        s = torch.cuda.Stream(dev)
        with torch.cuda.stream(s):
            path()
            if par is not None:
                exchange()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            path()
            if par is not None:                     # fixed launch parameters: the exchange sits in the step's graph
                exchange()

    def step():
        if graph is not None:
            graph.replay()
            if par is None:
                exchange()
        else:
            path()
            exchange()

    def timed(fn, k, w):
        for _ in range(w):
            l2_flush(); coll.rank_sync(); fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(k)]
        for a, b in ev:
            l2_flush()
            coll.rank_sync()
            a.record(); fn(); b.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        tot = sum(a.elapsed_time(b) for a, b in ev)
        if world > 1:
            t = torch.tensor([tot], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            tot = float(t.item())
        return tot                                   # ms for `k` steps

    sampler = ClockSampler(dev.index)
    sampler.start()
    tot_ms = timed(step, steps, warmup)
    clocks = sampler.stop()
    ms_per_step = tot_ms / steps

    # the all-reduced buffer of the LAST timed step against the closed form, bit for bit (outside the timed region)
    coll_check = "n/a"
    if world > 1:
        okv = torch.tensor([1.0 if torch.equal(psum, pgrad_expected(npar, world, dev)) else 0.0], device=dev)
        dist.all_reduce(okv, op=dist.ReduceOp.MIN)
        coll_check = "ok" if okv.item() >= 1.0 else "mismatch"

    vn = out.voxel_num.cpu().numpy()
    M = int(vn[-1])
    K = int((out.point2voxel >= 0).sum().item())
    launches_per_step = 2 + (1 if par is not None else 0)   # forward (persistent, cooperative) + k_step_bwd [+ all-reduce]

    # ---- e2e: pinned host buffers in, every API-visible output back to pinned host buffers, through the
    # package's public host front end (StepPipeline: H2D, kernels and D2H on three streams, double-buffered,
    # every step's copies inside the timed region); the perturber-gradient all-reduce follows each submit.
    pipe = StepPipeline(cfg, bounds, n, dev, depth=2, dynamic=dynamic, use_graph=not args.no_graph, pre_step=l2_flush)
    h_in_pts, h_in_raw = pipe.alloc_host_inputs()           # the caller's pinned staging buffer (one copy per step)
    h_in_pts.copy_(h_pts)
    h_in_raw.copy_(h_raw)
    e2e_steps = max(3, min(steps, 100))

    def e2e_measure(pipe, with_raw=True):
        def e2e_run(k):
            for _ in range(k):
                pipe.submit(h_in_pts, h_in_raw if with_raw else None, gfe)
                if world > 1 and par is None:
                    psum.copy_(pgrad)
                    dist.all_reduce(psum)
                elif par is not None:
                    with torch.cuda.stream(pipe.s_cmp):
                        par.all_reduce(pgrad, psum)
            pipe.drain()

        e2e_run(max(3, min(warmup, 5)))
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for s3 in (pipe.s_in, pipe.s_cmp, pipe.s_out):
            s3.wait_event(ev0)
        e2e_run(e2e_steps)
        for s3 in (pipe.s_in, pipe.s_cmp, pipe.s_out):
            torch.cuda.current_stream().wait_stream(s3)
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # consecutive blocks of e2e_steps; the MEDIAN block is reported and all are listed (on this pool one block in a few
    # comes out with the two PCIe directions not overlapping -- the VM's copy-engine scheduling, not the pipeline)
    nblocks = 5 if full else 3
    e2e_blocks = [e2e_measure(pipe) for _ in range(nblocks)]
    e2e_ms = float(np.median(e2e_blocks))
    h2d, d2h = pipe.h2d_bytes, pipe.d2h_bytes
    del pipe
    # the same call with host_fields=LEAN_FIELDS: the zero-padded voxels tensor (an intermediate once the VFE mean is
    # fused; the reference hands feats + coors on to the middle encoder) stays on the device.
    lean = None
    if full and not dynamic:
        lean_pipe = StepPipeline(cfg, bounds, n, dev, depth=2, dynamic=False, use_graph=not args.no_graph,
                                 pre_step=l2_flush, host_fields=StepPipeline.LEAN_FIELDS)
        lb = [e2e_measure(lean_pipe) for _ in range(3)]
        lean = (float(np.median(lb)), lean_pipe.d2h_bytes, lb)
        del lean_pipe

    # what the reference's training step would move and nothing else: points in; feats, coors,
    # num_points, voxel_num out; the perturbation is resident on the device (where the perturber produces it) and its
    # gradient stays there (where the perturber's backward consumes it).  The perturber's own cost is not in this number:
    # profiles/r02_mlp_in_loop.txt has the step with the real MLP.
    real = None
    if full and not dynamic:
        real_pipe = StepPipeline(cfg, bounds, n, dev, depth=2, dynamic=False, use_graph=not args.no_graph, pre_step=l2_flush,
                                 host_fields=StepPipeline.LEAN_FIELDS, raw_on_device=True, grad_to_host=False)
        real_pipe.set_raw(raw)
        rb = [e2e_measure(real_pipe, with_raw=False) for _ in range(3)]
        real = (float(np.median(rb)), real_pipe.h2d_bytes, real_pipe.d2h_bytes, rb)
        del real_pipe

    # ---- per-kernel durations (CUDA events inside the library, same steps, cold L2, no graph)
    stages, phases = {}, {}
    _lib.profile_enable(True)
    prof_steps = max(3, min(steps, 30))
    for it in range(prof_steps + 2):
        l2_flush()
        fwd()
        t_fwd = _lib.profile_read()
        ph = (ops.dyn_phase_times_us(cfg, n, workspace=ws) if dynamic else ops.phase_times_us(cfg, n, workspace=ws))
        bwd()
        t_bwd = _lib.profile_read()
        if it >= 2:
            for nm, ms in t_fwd + t_bwd:
                stages.setdefault(nm, []).append(ms)
            for nm, us in ph.items():
                phases.setdefault(nm, []).append(us)
    _lib.profile_enable(False)
    stage_us = {k2: round(1e3 * float(np.mean(v)), 3) for k2, v in stages.items()}
    peak, peak_src = measured_peak_gbs()
    balg = b_alg_dyn(N, C, M, F) if dynamic else b_alg_hard(N, C, P, M, K, F)
    # this repo's own per-kernel byte model (DESIGN.md "kernels"): what that kernel alone must read and write
    kern_bytes = {
        "k_hard_fwd": 4 * (2 * N * C + M * P * C + 5 * M + M * F + N),
        "k_dyn_scatter": 4 * (2 * N * C + 5 * N + 5 * M + M * F),
        "k_step_bwd": 4 * (N + N * C + M * F + M + N * C),
    }
    dom = max(stage_us, key=stage_us.get) if stage_us else None
    step_ach = balg / (ms_per_step * 1e-3) / 1e9
    # roofline.frac: algorithmic bytes of the WHOLE step / duration of the dominant kernel / measured peak;
    # step_*: the same bytes over the whole graph-replayed step (what `value` is computed from)
    roof = {"bound": "hbm", "kernel": dom, "unit": "GB/s", "peak": peak, "peak_source": peak_src, "traffic": None,
            "b_alg_bytes": balg, "stage_us": stage_us,
            "fwd_phase_us": {k2: round(float(np.mean(v)), 2) for k2, v in phases.items()},
            "step_b_alg_bytes": balg, "step_achieved": round(step_ach, 2), "step_frac": round(step_ach / peak, 4),
            "step": {"b_alg_bytes": balg, "achieved": round(step_ach, 2), "frac": round(step_ach / peak, 4)}}
    if dom is not None:
        if name == args.workload:
            roof.update(ncu_traffic(args, dom))
        a = balg / (stage_us[dom] * 1e-6) / 1e9
        roof.update({"achieved": round(a, 2), "frac": round(a / peak, 4),
                     "kernel_bytes_model": {k2: kern_bytes[k2] for k2 in stage_us if k2 in kern_bytes}})
    else:
        roof.update({"achieved": None, "frac": None})
    total_pts = N * world
    res = {
        "value": round(total_pts / (ms_per_step * 1e-3) / 1e6, 2), "ms_per_step": round(ms_per_step, 5),
        "config": config_dict(name, p, B, l2="flushed between steps, outside the per-step CUDA-event pairs: 512 MiB overwrite" +
                              (" then a 256 MiB read-only sweep (L2 left full of clean lines)"
                               if args.flush == "write+read" else ""), cuda_graph=graph is not None, voxels=M, kept_points=K,
                              collective=coll.describe(npar), collective_check=coll_check),
        "clocks": clocks,
        "e2e": {"value": round(total_pts / (e2e_ms / e2e_steps * 1e-3) / 1e6, 2), "unit": "M points/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                "ms_per_step": round(e2e_ms / e2e_steps, 5),
                "blocks_ms_per_step": [round(b / e2e_steps, 5) for b in e2e_blocks], "block_stat": "median",
                "how": "StepPipeline: pinned host in/out, ONE H2D and ONE D2H copy per step (packed buffers), H2D | kernels | D2H on three streams, depth 2"},
        "gpu_launches": launches_per_step * steps,
        "roofline": roof,
    }
    if lean is not None:
        res["e2e_lean"] = {"value": round(total_pts / (lean[0] / e2e_steps * 1e-3) / 1e6, 2), "unit": "M points/s",
                           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": lean[1],
                           "ms_per_step": round(lean[0] / e2e_steps, 5),
                           "blocks_ms_per_step": [round(b / e2e_steps, 5) for b in lean[2]],
                           "how": "as e2e, StepPipeline(host_fields=LEAN_FIELDS): feats, coors, num_points, voxel_num and grad_raw return to the host, the zero-padded voxels stay on the device"}
    if real is not None:
        res["e2e_real"] = {"value": round(total_pts / (real[0] / e2e_steps * 1e-3) / 1e6, 2), "unit": "M points/s",
                           "h2d_bytes_per_step": real[1], "d2h_bytes_per_step": real[2],
                           "ms_per_step": round(real[0] / e2e_steps, 5),
                           "blocks_ms_per_step": [round(b / e2e_steps, 5) for b in real[3]],
                           "how": "StepPipeline(raw_on_device=True, grad_to_host=False, host_fields=LEAN_FIELDS): points in; feats, coors, num_points, voxel_num out; the perturbation and its gradient stay on the device (stand-in perturbation, the MLP's own cost is in profiles/r02_mlp_in_loop.txt)"}
    if full and world == 1 and not args.no_cpu:
        res["cpu_baseline"] = cpu_baseline(name, clouds, raws, p)
    return res


def run_b200(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    from robustpointclouds_b200.hostmem import bind_to_gpu_numa
    numa_cpus = None if args.no_numa else bind_to_gpu_numa(local)      # pinned host buffers on the GPU's own NUMA node
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    coll = Collective(args, world, rank, dev)

    head = measure(args, args.workload, args.batch, args.points, args.steps, args.warmup, coll, rank, world, dev, True)
    # BASELINE.json configs[2], [3] and the top of the configs[4] sweep in the same invocation (shorter runs); at 8 GPUs
    # the dynamic workload with ONE sweep per rank is configs[3] as written ("batch-sharded on 8 x B200")
    also = {}
    if not args.no_also and not args.batch and not args.points and args.workload == "kitti":
        k = max(10, min(args.steps, 60))
        extra = [("nuscenes", "nuscenes", 0, 0), ("nuscenes-dynamic", "nuscenes-dynamic", 0, 0)]
        if world == 8:
            extra.append(("nuscenes-dynamic-1-sweep-per-gpu", "nuscenes-dynamic", 1, 0))
        if world == 1:
            extra.append(("kitti-2x2M", "kitti", 2, 2_000_000))
        for key, wl, b, pts_n in extra:
            r = measure(args, wl, b, pts_n, k, max(3, min(args.warmup, 5)), coll, rank, world, dev, False)
            also[key] = {"value": r["value"], "unit": "M points/s", "ms_per_step": r["ms_per_step"], "steps": k,
                         "workload": r["config"]["workload"], "voxels": r["config"]["voxels"],
                         "collective_check": r["config"]["collective_check"],
                         "roofline": {q: r["roofline"][q] for q in ("kernel", "b_alg_bytes", "achieved", "frac", "stage_us",
                                                                   "step_achieved", "step_frac")},
                         "e2e": {q: r["e2e"][q] for q in ("value", "ms_per_step", "h2d_bytes_per_step",
                                                          "d2h_bytes_per_step", "blocks_ms_per_step")}}
    ar_timeouts = coll.timeouts()
    if ar_timeouts:
        print(f"[bench] rank {rank}: peer all-reduce saw a timeout", file=sys.stderr)
    if rank == 0:
        head["config"]["collective_timeouts"] = ar_timeouts
        if numa_cpus:
            head["e2e"]["how"] += "; process bound to the %d CPUs local to the GPU" % len(numa_cpus)
        line = {
            "metric": "M points/s perturb+voxelize+VFE fwd/bwd", "value": head["value"],
            "unit": "M points/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": head["config"], "clocks": head["clocks"],
            "e2e": head["e2e"], "e2e_lean": head.get("e2e_lean"), "e2e_real": head.get("e2e_real"), "gpu_launches": head["gpu_launches"],
            "roofline": head["roofline"],
        }
        if "cpu_baseline" in head:
            line["cpu_baseline"] = head["cpu_baseline"]
        if also:
            line["also"] = also
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="kitti", choices=["kitti", "nuscenes", "nuscenes-dynamic", "kitti-dynamic"])
    ap.add_argument("--batch", type=int, default=0, help="sweeps per GPU (default: 6 KITTI, 4 nuScenes)")
    ap.add_argument("--points", type=int, default=0, help="points per sweep (default: the preset's; BASELINE.json configs[4] sweep)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-numa", action="store_true", help="do not bind the process to the GPU's local CPUs")
    ap.add_argument("--allreduce", default="ll", choices=["ll", "p2p", "nccl"],
                    help="N > 1: perturber-gradient all-reduce through this library's one-shot NVLink kernel or NCCL")
    ap.add_argument("--flush", default="write", choices=["write", "write+read"],
                    help="L2 flush between steps: 512 MiB overwrite, optionally followed by a 256 MiB read sweep")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-also", action="store_true", help="headline workload only (no nuscenes / dynamic / 2 x 2M entries)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    global POINTS_OVERRIDE
    POINTS_OVERRIDE = args.points
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
