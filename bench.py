#!/usr/bin/env python
"""bench.py - sample+group points/sec (BASELINE.json metric) on B200.

One step = one pass of the hot path over the batch of synthetic tooth-scan-shaped clouds BASELINE.json names:
    furthest_point_sample(M=1024) -> gather centres -> ball_query(r=0.1, ns=32)
    -> grouping_operation(xyz, C=3) -> grouping_operation(features, C=64)
on the GLOBAL batch of B=64 clouds x N=16384 points.  With N GPUs (torchrun, one rank per GPU) the batch
is sharded by cloud - 64/N clouds per rank, no data-path collective - and every step ends with the one exchange the path
has: the NCCL all-gather / all-reduce of the transition-matrix class statistics and T-predictor gradients
(geot_b200/dist.py), issued on a side stream (`scaling: "strong"`).
`value` = 64 x 16384 points / the step time with inputs resident in HBM: ONE step in flight on one pre-warmed stream,
CUDA events around every step, median over the K timed steps, max over ranks.  `kernels` / `roofline` come
from CUDA events between the ops of those same timed steps.  `e2e` = the same pass with HOST buffers: pinned host ->
device copy of xyz + features every step and a device -> host read of the sampled / neighbour indices and centres
inside the timed region.  `run` holds what is not the headline: weak scaling (64 clouds per GPU), the B=256 batch, and
the throughput with three different batches in flight.

    python bench.py [--gpus N --steps K --warmup W]          our arm
    python bench.py --impl reference ...                       the reference semantics on the host cores
    python bench.py --config cfg2|cfg3|cfg4                    BASELINE configs 2-4: reference models, both backends
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

B, N, M, NS, C, RADIUS = 64, 16384, 1024, 32, 64, 0.1
C_STAT = 33  # classes of the transition-matrix statistics exchanged per step when n_gpus > 1 (BASELINE config 4)
SPIN_UP_STEPS = 200  # untimed steps before the warm-up (see run_ours)
WORKLOAD = f"sample+group B={B}xN={N}: FPS(M={M}) + gather + ball_query(r={RADIUS},ns={NS}) + grouping(C=3) + grouping(C={C})"
METRIC = "sample+group points/sec at B=64,N=16k"


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# Algorithmic HBM bytes per launch, for B clouds.
def algo_bytes(b=B):
    return {
        "fps": b * (12 * N + 4 * M),
        "gather_centres": b * (4 * M + 2 * 4 * 3 * M),
        "ball_query": b * (12 * N + 12 * M + 4 * M * NS),
        "group_xyz": b * (4 * M * NS + 4 * 3 * N + 4 * 3 * M * NS),
        "group_feat": b * (4 * M * NS + 4 * C * N + 4 * C * M * NS),
    }


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt, self.proc = index, [], threading.Event(), None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
                if self._stop_evt.is_set():
                    break
        except Exception:
            pass

    def stop(self):
        self._stop_evt.set()
        if self.proc is not None:
            try:
                self.proc.terminate()
            except Exception:
                pass

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def cpu_pass(xyz_h, feat_h, oracle):
    """One pass of the same workload with the CPU oracle (the reference semantics, all host threads)."""
    idx = oracle.furthest_point_sample(xyz_h, M)
    centres = np.take_along_axis(xyz_h, idx[..., None].astype(np.int64).repeat(3, 2), 1)
    nbr = oracle.ball_query(RADIUS, NS, xyz_h, centres)
    gx = oracle.grouping_operation(np.ascontiguousarray(xyz_h.transpose(0, 2, 1)), nbr)
    gf = oracle.grouping_operation(feat_h, nbr)
    return idx, nbr, gx, gf


def run_reference(args):
    """--impl reference: the reference has NO CPU implementation of these ops (pointnet2/_ext_src/src/
    ball_query.cpp:31 'CPU not supported'), so this arm times the CPU port of its kernel semantics
    (oracle/cpu_ref.c, OpenMP over all host threads) on the same workload.  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from geot_b200 import synth
    from oracle import cpu_ref as oracle
    oracle.load()
    oracle.set_num_threads(os.cpu_count() or 1)  # torchrun exports OMP_NUM_THREADS=1
    xyz_h = synth.tooth_batch(B, N)
    feat_h = np.random.default_rng(0).standard_normal((B, C, N)).astype(np.float32)
    for _ in range(args.warmup):
        cpu_pass(xyz_h, feat_h, oracle)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_pass(xyz_h, feat_h, oracle)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    val = B * N / dt
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "points/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": bench_config(int(os.environ.get("WORLD_SIZE", "1"))),
            "cpu_baseline": {"value": val, "unit": "points/s", "cores": oracle.num_threads(), "kind": "port",
                             "sample": f"the full B={B}xN={N} batch per step (one step = {dt:.2f} s of wall time on {oracle.num_threads()} threads)"},
            "e2e": {"value": val, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def bind_to_gpu_cpus(local):
    """Multi-rank runs: pin this process (and so the first touch of its pinned host buffers) to the CPUs NVML reports
    as local to its GPU, so that the per-step host->device copies of the ranks do not all cross one socket.
    Returns the number of CPUs bound to (0 = left as is)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        allowed = os.sched_getaffinity(0)
        cpus = [64 * i + b for i, w in enumerate(mask) for b in range(64) if (w >> b) & 1 and (64 * i + b) in allowed]
        if cpus and len(cpus) < len(allowed):
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return 0


def bench_config(world):
    """The `config` object - identical in both arms (ours / --impl reference) for the same n_gpus."""
    per = B // world
    foot = per * (12 * N + 4 * C * N + 4 * M + 12 * M + 4 * M * NS + 12 * M * NS + 4 * C * M * NS) / 1e6
    return {"workload": WORKLOAD, "global_clouds": B, "points_per_cloud": N, "clouds_per_gpu": per,
            "sharding": "strong scaling: the global B=64 batch is sharded by cloud (no data-path collective); with n_gpus > 1 every "
                        f"step ends with the NCCL exchange of the transition-matrix class statistics and T-predictor gradients (C={C_STAT})",
            "l2_policy": f"no flush: a step touches {foot:.0f} MB per GPU (inputs + outputs) and consecutive steps rotate over "
                         f"{n_rotating_sets(world)} disjoint buffer sets (> 3x the 126 MB L2 between two uses of an address)",
            "timing": "CUDA events per step on one pre-warmed stream (inside a step the ball query's binning of the clouds runs on a side stream while FPS computes the centres and is joined before the queries); ms_per_step = median over the K timed steps, max over ranks; 200 untimed spin-up steps precede the W warm-up steps"}


def n_rotating_sets(world):
    per = B // world
    foot = per * (12 * N + 4 * C * N + 4 * M + 12 * M + 4 * M * NS + 12 * M * NS + 4 * C * M * NS)
    return max(1, -(-400_000_000 // foot))


class Workset:
    """Device buffers of one batch of `b` clouds and the five ops of the hot path over them."""

    def __init__(self, torch, F, _lib, dev, b, xyz_h=None, feat_h=None, first_cloud=0, seed=0):
        from geot_b200 import synth
        i32, f32 = torch.int32, torch.float32
        self.b = b
        self.xyz_h = xyz_h if xyz_h is not None else torch.from_numpy(synth.tooth_batch(b, N, first_cloud=first_cloud)).pin_memory()
        self.feat_h = feat_h if feat_h is not None else torch.from_numpy(
            np.random.default_rng(seed).standard_normal((b, C, N)).astype(np.float32)).pin_memory()
        self.xyz, self.feat = self.xyz_h.to(dev), self.feat_h.to(dev)
        self.fps_idx = torch.empty((b, M), dtype=i32, device=dev)
        self.centres = torch.empty((b, M, 3), dtype=f32, device=dev)
        self.nbr = torch.empty((b, M, NS), dtype=i32, device=dev)
        self.g_xyz = torch.empty((b, 3, M, NS), dtype=f32, device=dev)
        self.g_feat = torch.empty((b, C, M, NS), dtype=f32, device=dev)
        self.bq_ws = torch.empty(int(_lib.lib().geot_ball_query_workspace_bytes(b, N)), dtype=torch.uint8, device=dev)
        s = self
        # the ball query's binning of the clouds needs xyz, the radius and nsample, not the centres: it runs on a side
        # stream while FPS (one CTA per cloud) computes them and is joined before the queries (F.fps_ball_query is the
        # same sequence as one call; here the ops stay separate for the per-op events)
        self.side = torch.cuda.Stream(device=dev)
        self.ev_ready, self.ev_built = torch.cuda.Event(), torch.cuda.Event()

        def fps_with_binning():
            cur = torch.cuda.current_stream()
            s.ev_ready.record(cur)
            s.side.wait_event(s.ev_ready)
            with torch.cuda.stream(s.side):
                F.ball_query_prepare(s.xyz, RADIUS, NS, s.bq_ws)
                s.ev_built.record(s.side)
            F.furthest_point_sampling(s.xyz, M, out=s.fps_idx)

        def ball_query_prebuilt():
            torch.cuda.current_stream().wait_event(s.ev_built)
            F.ball_query(s.centres, s.xyz, RADIUS, NS, out=s.nbr, workspace=s.bq_ws, flags=_lib.GEOT_BQ_PREBUILT)

        self.ops = {
            "fps": fps_with_binning,
            # centres = xyz[fps_idx], gathered point-major: same values as the reference's transpose -> gather_operation
            # -> transpose chain (checked once per run), without the two copies
            "gather_centres": lambda: F.gather_rows(s.xyz, s.fps_idx, out=s.centres),
            "ball_query": ball_query_prebuilt,
            # grouping_operation(xyz^T, idx) straight from the (B,N,3) cloud (no (B,3,N) copy): identical output
            "group_xyz": lambda: F.group_xyz_relative(s.xyz, s.centres, s.nbr, relative=False, out=s.g_xyz),
            "group_feat": lambda: F.group_points(s.feat, s.nbr, out=s.g_feat),
        }

    def step(self):
        for f in self.ops.values():
            f()

    def check_against_composite(self, torch, F):
        """One-time check that the two fused ops equal the reference's composite chain on this batch."""
        self.step()
        xyz_t = self.xyz.transpose(1, 2).contiguous()
        centres_t = F.gather_points(xyz_t, self.fps_idx)
        assert torch.equal(self.centres, centres_t.transpose(1, 2)), "gather_rows != transpose/gather_operation/transpose"
        assert torch.equal(self.g_xyz, F.group_points(xyz_t, self.nbr)), "group_xyz_relative(relative=False) != grouping_operation(xyz^T)"


class StatsExchange:
    """The one collective of the path (geot_b200/dist.py StatsExchange): per step, on a side stream, ONE
    packed all-gather of the class-anchor rows (C x (C+1) floats), the count-weighted class means (C*C + C) and the
    T-predictor gradients (C*C*2C floats = 287 KB at C=33), combined by one kernel in fixed rank order.  The statistics
    themselves are computed once, outside the timed region, by the transition kernels on this rank's shard of a
    synthetic teacher output."""

    def __init__(self, torch, dev, rank, world, clouds):
        from geot_b200 import dist as gdist
        from geot_b200 import transition as gt
        self.world = world
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        pred = torch.softmax(torch.randn((max(clouds, 1), C_STAT, 2048), device=dev, generator=g) * 2, 1)
        self.class_T, self.best_v, _ = gt.class_T_pick(pred)
        self.cm = torch.softmax(torch.randn((C_STAT, C_STAT), device=dev, generator=g), 1)
        self.c_num = torch.rand((C_STAT,), device=dev, generator=g) * 100
        self.params = [torch.nn.Parameter(torch.zeros((C_STAT, 2 * C_STAT), device=dev)) for _ in range(C_STAT)]
        for p_ in self.params:
            p_.grad = torch.randn(p_.shape, device=dev, generator=g)
        self.ex = gdist.StatsExchange(C_STAT, self.params)
        self.bytes = 4 * (2 * C_STAT * C_STAT + 2 * C_STAT + C_STAT * C_STAT * 2 * C_STAT)
        self.result = None

    def launch(self):
        """Enqueue the exchange on its side stream (it does not depend on this step's sampling / grouping)."""
        if self.world > 1:
            self.ex.launch(self.class_T, self.best_v, self.cm, self.c_num)

    def join(self):
        if self.world > 1:
            self.result = self.ex.wait()


def run_ours(args):
    import torch
    import torch.distributed as dist
    from geot_b200 import _lib
    from geot_b200 import dist as gdist
    from geot_b200 import functional as F

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback exists)"
    assert B % world == 0, "the global batch of 64 clouds is sharded evenly: use 1, 2, 4 or 8 GPUs"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_cpus = bind_to_gpu_cpus(local) if world > 1 else 0  # before the pinned buffers are first touched
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner to stdout on the first communicator; stdout must carry ONE JSON line
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def rank_max(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    # ---- the measured configuration: the GLOBAL batch of 64 clouds, sharded by cloud over the ranks (strong scaling)
    first, per = gdist.shard_clouds(B, rank, world)
    nsets = n_rotating_sets(world)
    sets = [Workset(torch, F, _lib, dev, per, first_cloud=s * B + first, seed=1000 * s + rank) for s in range(nsets)]
    for ws in sets:
        ws.check_against_composite(torch, F)
    stats = StatsExchange(torch, dev, rank, world, per)
    op_names = list(sets[0].ops)

    def timed_steps(steps, warmup, fn):
        """CUDA events around every one of exactly `steps` calls of fn(i) on the current (pre-warmed) stream.
        Returns the per-step ms list of this rank."""
        for i in range(warmup):
            fn(i)
        barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        ev[0].record()
        for i in range(steps):
            fn(warmup + i)
            ev[i + 1].record()
        barrier()
        return [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]

    # main timed region: one step in flight, per-op events inside every step (the shares below come from the SAME region)
    op_ev = []

    def step_main(i, record=True):
        ws = sets[i % nsets]
        stats.launch()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(len(op_names) + 1)] if record else None
        if record:
            evs[0].record()
        for k, f in enumerate(ws.ops.values()):
            f()
            if record:
                evs[k + 1].record()
        stats.join()
        if record:
            op_ev.append(evs)

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    # untimed spin-up before the W warm-up steps: a GPU that has just left idle needs some 0.1 s under load before its
    # clocks settle (seen on the second GPU of a fresh box: FPS 0.64 ms for the first steps against 0.56 ms).  A fixed
    # count, so that every rank makes the same number of exchange calls.
    for i in range(SPIN_UP_STEPS):
        step_main(i, record=False)
    barrier()
    l0 = _lib.launch_count()
    per_step = timed_steps(args.steps, args.warmup, step_main)
    launches_per_step = (_lib.launch_count() - l0) // (args.steps + args.warmup)
    op_ev = op_ev[args.warmup:]
    ms_med = rank_max(float(np.median(per_step)))
    ms_mean = rank_max(float(np.mean(per_step)))
    per_op = {n_: rank_max(float(np.median([e[k].elapsed_time(e[k + 1]) for e in op_ev]))) for k, n_ in enumerate(op_names)}

    # e2e: HOST buffers every step.  Two device input sets; the host->device copy of step i+1 runs on a copy stream
    # while step i computes, results (sampled indices, neighbour indices, centres) are read back on the compute stream.
    ws0 = sets[0]
    e2e_ws = [ws0, Workset(torch, F, _lib, dev, per, xyz_h=ws0.xyz_h, feat_h=ws0.feat_h)]
    out_h = [torch.empty((per, M), dtype=torch.int32).pin_memory(), torch.empty((per, M, NS), dtype=torch.int32).pin_memory(),
             torch.empty((per, M, 3), dtype=torch.float32).pin_memory()]
    copy_stream = torch.cuda.Stream(device=dev)
    ev_copied = [torch.cuda.Event(), torch.cuda.Event()]
    ev_free = [torch.cuda.Event(), torch.cuda.Event()]

    def step_e2e(i):
        k = i & 1
        ws = e2e_ws[k]
        cur = torch.cuda.current_stream()
        stats.launch()
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(ev_free[k])  # the step that last used this input set has finished
            ws.xyz.copy_(ws.xyz_h, non_blocking=True)
            ws.feat.copy_(ws.feat_h, non_blocking=True)
            ev_copied[k].record(copy_stream)
        cur.wait_event(ev_copied[k])
        ws.step()
        ev_free[k].record(cur)
        out_h[0].copy_(ws.fps_idx, non_blocking=True)
        out_h[1].copy_(ws.nbr, non_blocking=True)
        out_h[2].copy_(ws.centres, non_blocking=True)
        stats.join()

    e2e_steps = timed_steps(args.steps, max(args.warmup, 3), step_e2e)
    ms_e2e = rank_max(float(np.sum(e2e_steps)) / args.steps)  # pipelined copies: the mean over the region is the step cost
    if sampler:
        sampler.stop()

    # ---- extra lines (not the headline): weak scaling (64 clouds on EVERY GPU), the B=256 global batch, and the
    # throughput with three independent batches in flight on three streams
    extra_steps = min(args.steps, 20)
    extras = {}
    if not args.no_extras:
        del e2e_ws
        def serial_of(wsets):
            it = timed_steps(extra_steps, 3, lambda i: (stats.launch(), wsets[i % len(wsets)].step(), stats.join()))
            return rank_max(float(np.median(it)))
        if world > 1:
            wsets = [Workset(torch, F, _lib, dev, B, first_cloud=rank * B, seed=rank)]
            ms_w = serial_of(wsets)
            extras["weak_64_clouds_per_gpu"] = {"ms_per_step": ms_w, "points_per_s": B * N * world / (ms_w * 1e-3), "scaling": "weak"}
            del wsets
        b256 = 256 // world
        wsets = [Workset(torch, F, _lib, dev, b256, first_cloud=5000 + rank * b256, seed=50 + rank)]
        ms_256 = serial_of(wsets)
        extras["global_batch_256"] = {"clouds_per_gpu": b256, "ms_per_step": ms_256, "points_per_s": 256 * N / (ms_256 * 1e-3)}
        del wsets
        if world == 1:
            # three DIFFERENT batches in flight on three streams (a second / third batch must exist for this to be real)
            psets = [sets[0]] + [Workset(torch, F, _lib, dev, per, first_cloud=7000 + 64 * k, seed=70 + k) for k in range(2)]
            streams = [torch.cuda.Stream(device=dev) for _ in psets]
            for _ in range(4):
                for st, ws in zip(streams, psets):
                    with torch.cuda.stream(st):
                        ws.step()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            cur = torch.cuda.current_stream()
            e0.record()
            for st in streams:
                st.wait_event(e0)
            for i in range(extra_steps):
                k = i % len(psets)
                with torch.cuda.stream(streams[k]):
                    psets[k].step()
            for st in streams:
                cur.wait_stream(st)
            e1.record()
            barrier()
            ms_o = e0.elapsed_time(e1) / extra_steps
            extras["overlapped_3_batches_in_flight"] = {"ms_per_step": ms_o, "points_per_s": B * N / (ms_o * 1e-3),
                                                        "note": "three different batches on three streams; throughput, not step latency"}

    line = None
    if rank == 0:
        peak, peak_src = peaks()
        ab = algo_bytes(per)
        ksum = sum(per_op.values())
        kern = {k: {"ms": per_op[k], "share": per_op[k] / ksum, "algo_bytes": ab[k], "hbm_gbs": ab[k] / per_op[k] / 1e6,
                    "hbm_frac": ab[k] / per_op[k] / 1e6 / peak} for k in per_op}
        # FPS is not HBM-bound: M-1 dependent rounds; report its own latency view too
        kern["fps"].update({"bound": "dependent-latency chain (profiles/r02_fps_study.md)", "us_per_round": per_op["fps"] * 1e3 / (M - 1),
                            "clouds_in_flight": per})
        gf = kern["group_feat"]
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("group_feat_dram_bytes")
            traffic = traffic * per // B if traffic else None
        except Exception:
            pass
        total_pts = B * N
        line = {"metric": METRIC, "value": total_pts / (ms_med * 1e-3), "unit": "points/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_med, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": bench_config(world),
                "run": {"ms_per_step_mean": ms_mean, "steps_in_flight": 1, "rotating_buffer_sets": nsets, "rank_cpu_binding": numa_cpus,
                        "stats_exchange_bytes_per_step": stats.bytes if world > 1 else 0, **extras},
                "e2e": {"value": total_pts / (ms_e2e * 1e-3), "unit": "points/s", "ms_per_step": ms_e2e,
                        "h2d_bytes_per_step": (ws0.xyz_h.numel() + ws0.feat_h.numel()) * 4 * world,
                        "d2h_bytes_per_step": sum(t.numel() for t in out_h) * 4 * world,
                        "note": "whole job (all ranks): every step copies xyz + features from pinned host memory and reads back the sampled "
                                "indices, the neighbour indices and the centres; the grouped tensors (C=3 and C=64, 562 MB per 64 clouds) stay "
                                "on the device for the consumer (the shared MLP), they are NOT copied back"},
                "gpu_launches": int(launches_per_step * args.steps * world),
                "roofline": {"kernel": "group_points (features, C=64): the dominant HBM-bound kernel; FPS dominates time but is latency-bound, see kernels",
                             "bound": "hbm", "achieved": gf["hbm_gbs"], "peak": peak, "unit": "GB/s", "frac": gf["hbm_frac"],
                             "traffic": traffic, "peak_source": peak_src,
                             "how": "algorithmic bytes of one launch / median CUDA-event time of that launch inside the timed steps"},
                "kernels": kern, "clocks": sampler.summary() if sampler else None}
    if world > 1:
        dist.barrier()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import cpu_ref as oracle  # checker/baseline leg only; never on the product path
        oracle.load()
        oracle.set_num_threads(os.cpu_count() or 1)
        xs, fs = ws0.xyz_h.numpy(), ws0.feat_h.numpy()
        best = 1e30
        for _ in range(3):
            t0 = time.perf_counter()
            cpu_pass(xs, fs, oracle)
            best = min(best, time.perf_counter() - t0)
        line["cpu_baseline"] = {"value": B * N / best, "unit": "points/s", "cores": oracle.num_threads(), "kind": "port",
                                "sample": f"the full B={B}xN={N} batch, best of 3 passes ({best:.2f} s each), oracle/cpu_ref.c OpenMP"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_config(args):
    """--config cfg2|cfg3|cfg4: BASELINE.json configs 2-4 = the reference's OWN models / op wrappers (unchanged,
    tests/model_cases.py) timed on two backends on this GPU: geot_b200 through the pybind-shaped facades (scatter-plan
    builds included - the facades rebuild the plan per backward call) and the reference's own kernels
    (oracle/_ref/*.so).  Prints one JSON line: forward and forward+backward ms per backend (CUDA events, median) and
    the per-entry-point device time inside the facades.  Not the headline metric; a profiles/ artefact."""
    import torch
    from tests import model_cases as mc
    from tests import refenv
    torch.cuda.set_device(0)
    refenv.install("geot")
    name = args.config
    if name == "cfg3":
        model, data = mc.cfg3_model(), mc.cfg3_inputs(16, 16384)
        run = lambda be, bwd: mc.run_cfg3(be, model, data, None, backward=bwd)
        what = "PointNet++ MSG segmentation (SURVEY Appendix B), B=16 x 16384, fwd / fwd+bwd"
    elif name == "cfg4":
        cfg, model, _ = mc.cfg4_models(17)
        data, data_u = mc.cfg4_inputs(16000)
        run = lambda be, bwd: mc.run_cfg4_model(be, model, data, data_u, None, backward=bwd)
        what = "GeoT student WholePartSeg(PointTransformer_seg_T), 2 labeled + 2 strong + 2 weak clouds x 16000, fwd / fwd+bwd"
    else:
        inp, weights = mc.cfg2_inputs(8, 16384), mc.cfg2_weights()
        run = lambda be, bwd: mc.run_cfg2(be, inp, weights, backward=bwd)
        what = "Point-Transformer op pyramid (SURVEY Appendix C), offset format, B=8 x 16384, k=16, fwd / fwd+bwd"

    def timed(be, bwd, reps):
        ts = []
        for i in range(reps + 2):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run(be, bwd)
            e1.record()
            torch.cuda.synchronize()
            if i >= 2:
                ts.append(e0.elapsed_time(e1))
        return float(np.median(ts))

    out = {"config": name, "workload": what, "reps": args.steps, "backends": {}}
    for be in ("geot", "ref"):
        r = {"fwd_ms": timed(be, False, args.steps), "fwd_bwd_ms": timed(be, True, args.steps)}
        with refenv.timing() as tm:
            run(be, True)
        r["facade_ms"] = {k: {"calls": len(v), "ms": round(sum(v), 4)} for k, v in sorted(tm.items(), key=lambda kv: -sum(kv[1]))}
        r["facade_total_ms"] = round(sum(sum(v) for v in tm.values()), 4)
        out["backends"][be] = r
    g, r = out["backends"]["geot"], out["backends"]["ref"]
    out["speedup"] = {"fwd": r["fwd_ms"] / g["fwd_ms"], "fwd_bwd": r["fwd_bwd_ms"] / g["fwd_bwd_ms"],
                      "ops_only": r["facade_total_ms"] / g["facade_total_ms"]}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the weak-scaling / B=256 / overlapped extra measurements")
    ap.add_argument("--config", default=None, choices=["cfg2", "cfg3", "cfg4"],
                    help="time BASELINE config 2/3/4 (reference models, both backends) instead of the headline metric")
    args = ap.parse_args()
    if args.config:
        args.steps = min(args.steps, 20)
        run_config(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
