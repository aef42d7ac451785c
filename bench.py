#!/usr/bin/env python
"""bench.py — sample+group points/sec (BASELINE.json metric) on B200.

One step = one pass of the hot path over one batch of synthetic tooth-scan-shaped clouds:
    furthest_point_sample(M=1024) -> gather centres -> ball_query(r=0.1, ns=32)
    -> grouping_operation(xyz, C=3) -> grouping_operation(features, C=64)
on B=64 clouds x N=16384 points PER GPU (weak scaling: every rank owns 64 different clouds,
no data-path collective).  `value` = points of all ranks / device time (inputs resident in HBM), with
`--in-flight` (default 3) independent steps enqueued on as many streams, each with its own output set: one step
alone leaves 84 of the 148 SMs idle while its FPS runs (one CTA per cloud, M-1 dependent rounds), and consecutive
steps work on independent batches; `config.ms_per_step_one_in_flight` is the strictly serial number;
`e2e` = the same pass with HOST buffers: pinned host -> device copy of xyz+features every step and a
device -> host read of the sampled / neighbour indices and centres inside the timed region.

    python bench.py [--gpus N --steps K --warmup W]          our arm
    python bench.py --impl reference ...                       the reference semantics on the host cores
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
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": WORKLOAD},
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


def run_ours(args):
    import torch
    import torch.distributed as dist
    from geot_b200 import _lib, synth
    from geot_b200 import functional as F

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback exists)"
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

    # every rank owns B different clouds (weak scaling, sharded by cloud, no data-path collective)
    xyz_h = torch.from_numpy(synth.tooth_batch(B, N, first_cloud=rank * B)).pin_memory()
    feat_h = torch.from_numpy(np.random.default_rng(rank).standard_normal((B, C, N)).astype(np.float32)).pin_memory()
    xyz = xyz_h.to(dev)
    feat = feat_h.to(dev)
    i32, f32 = torch.int32, torch.float32
    fps_idx = torch.empty((B, M), dtype=i32, device=dev)
    xyz_t = torch.empty((B, 3, N), dtype=f32, device=dev)
    centres_t = torch.empty((B, 3, M), dtype=f32, device=dev)
    centres = torch.empty((B, M, 3), dtype=f32, device=dev)
    nbr = torch.empty((B, M, NS), dtype=i32, device=dev)
    g_xyz = torch.empty((B, 3, M, NS), dtype=f32, device=dev)
    g_feat = torch.empty((B, C, M, NS), dtype=f32, device=dev)
    out_h = [torch.empty((B, M), dtype=i32).pin_memory(), torch.empty((B, M, NS), dtype=i32).pin_memory(),
             torch.empty((B, M, 3), dtype=f32).pin_memory()]
    bq_ws = torch.empty(int(_lib.lib().geot_ball_query_workspace_bytes(B, N)), dtype=torch.uint8, device=dev)

    def make_ops(xyz, feat):
        return {
            "fps": lambda: F.furthest_point_sampling(xyz, M, out=fps_idx),
            # centres = xyz[fps_idx], gathered point-major: same values as the reference's transpose -> gather_operation
            # -> transpose chain (checked once below), without the two copies
            "gather_centres": lambda: F.gather_rows(xyz, fps_idx, out=centres),
            "ball_query": lambda: F.ball_query(centres, xyz, RADIUS, NS, out=nbr, workspace=bq_ws),
            # grouping_operation(xyz^T, idx) straight from the (B,N,3) cloud (no (B,3,N) copy): identical output
            "group_xyz": lambda: F.group_xyz_relative(xyz, centres, nbr, relative=False, out=g_xyz),
            "group_feat": lambda: F.group_points(feat, nbr, out=g_feat),
        }

    ops = make_ops(xyz, feat)

    # `value`: several (default 3) steps in flight.  One step keeps 64 of the 148 SMs busy for its first 0.55 ms (FPS: one CTA per
    # cloud, M-1 dependent rounds); consecutive steps work on independent batches, so step i+1 is enqueued on a second
    # stream with its own output set and its FPS runs beside step i's ball query / grouping (and beside step i's FPS:
    # 2 x 64 CTAs fit).  Every step still runs all five ops on all 64 clouds inside the timed region.
    def out_set():
        return dict(fps_idx=torch.empty_like(fps_idx), centres=torch.empty_like(centres), nbr=torch.empty_like(nbr),
                    g_xyz=torch.empty_like(g_xyz), g_feat=torch.empty_like(g_feat), bq_ws=torch.empty_like(bq_ws))

    def make_ops_on(o):
        return [lambda: F.furthest_point_sampling(xyz, M, out=o["fps_idx"]),
                lambda: F.gather_rows(xyz, o["fps_idx"], out=o["centres"]),
                lambda: F.ball_query(o["centres"], xyz, RADIUS, NS, out=o["nbr"], workspace=o["bq_ws"]),
                lambda: F.group_xyz_relative(xyz, o["centres"], o["nbr"], relative=False, out=o["g_xyz"]),
                lambda: F.group_points(feat, o["nbr"], out=o["g_feat"])]

    NPIPE = max(1, args.in_flight)
    pipe_outs = [dict(fps_idx=fps_idx, centres=centres, nbr=nbr, g_xyz=g_xyz, g_feat=g_feat, bq_ws=bq_ws)] + [out_set() for _ in range(NPIPE - 1)]
    pipe_ops = [make_ops_on(o) for o in pipe_outs]
    pipe_streams = [torch.cuda.Stream(device=dev) for _ in range(NPIPE)]
    pipe_i = [0]

    def step_pipelined():
        k = pipe_i[0] % NPIPE
        pipe_i[0] += 1
        with torch.cuda.stream(pipe_streams[k]):
            for f in pipe_ops[k]:
                f()

    # one-time check that the two fused ops equal the reference's composite chain on this batch
    for f in ops.values():
        f()
    xyz_t.copy_(xyz.transpose(1, 2))
    F.gather_points(xyz_t, fps_idx, out=centres_t)
    assert torch.equal(centres, centres_t.transpose(1, 2)), "gather_rows != transpose/gather_operation/transpose"
    assert torch.equal(g_xyz, F.group_points(xyz_t, nbr)), "group_xyz_relative(relative=False) != grouping_operation(xyz^T)"

    def step():
        for f in ops.values():
            f()

    # e2e: HOST buffers every step.  Two device input sets; the host->device copy of step i+1 runs on a copy
    # stream while step i computes (the PCIe link, 281 MB per step, is the bound), results are read back on the
    # compute stream.  Every step's inputs cross the link inside the timed region.
    in_sets = [(xyz, feat), (torch.empty_like(xyz), torch.empty_like(feat))]
    ops_sets = [ops, make_ops(*in_sets[1])]
    copy_stream = torch.cuda.Stream(device=dev)
    ev_copied = [torch.cuda.Event(), torch.cuda.Event()]
    ev_free = [torch.cuda.Event(), torch.cuda.Event()]
    e2e_i = [0]

    def step_e2e():
        i = e2e_i[0] & 1
        e2e_i[0] += 1
        cur = torch.cuda.current_stream()
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(ev_free[i])  # the step that last used this input set has finished
            in_sets[i][0].copy_(xyz_h, non_blocking=True)
            in_sets[i][1].copy_(feat_h, non_blocking=True)
            ev_copied[i].record(copy_stream)
        cur.wait_event(ev_copied[i])
        for f in ops_sets[i].values():
            f()
        ev_free[i].record(cur)
        out_h[0].copy_(fps_idx, non_blocking=True)
        out_h[1].copy_(nbr, non_blocking=True)
        out_h[2].copy_(centres, non_blocking=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, side_streams=()):
        """CUDA events on the current stream around exactly `steps` calls; work that fn enqueues on side_streams starts
        after the first event and is waited for before the second."""
        for _ in range(warmup):
            fn()
        barrier()
        cur = torch.cuda.current_stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for st in side_streams:
            st.wait_event(e0)
        for _ in range(steps):
            fn()
        for st in side_streams:
            done = torch.cuda.Event()
            done.record(st)
            cur.wait_event(done)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    l0 = _lib.launch_count()
    ms_serial = timed(step, args.steps, args.warmup)
    launches = (_lib.launch_count() - l0) * world // (args.steps + args.warmup) * args.steps
    ms = timed(step_pipelined, args.steps, max(args.warmup, 4), side_streams=pipe_streams)
    # the second output set must hold what the first holds (same inputs): the pipelined steps did the full work
    assert all(torch.equal(o["nbr"], nbr) and torch.equal(o["g_feat"], g_feat) for o in pipe_outs[1:]), "pipelined step differs"
    ms_e2e = timed(step_e2e, args.steps, max(args.warmup, 3))
    if sampler:
        sampler.stop()

    # per-kernel device times (same stream, CUDA events), for the roofline block
    per = {}
    for name, fn in ops.items():
        per[name] = timed(fn, min(max(args.steps, 10), 50), 3)
    total_pts = B * N * world
    line = None
    if rank == 0:
        peak, peak_src = peaks()
        ab = algo_bytes()
        ksum = sum(per.values())
        kern = {k: {"ms": per[k], "share": per[k] / ksum, "algo_bytes": ab[k], "hbm_gbs": ab[k] / per[k] / 1e6,
                    "hbm_frac": ab[k] / per[k] / 1e6 / peak} for k in per}
        # FPS is not HBM-bound: M-1 sequential rounds of N distance updates; report its own issue-rate view too
        kern["fps"].update({"bound": "fp32-issue/latency", "rounds_per_s_per_cloud": (M - 1) / (per["fps"] * 1e-3),
                            "dist_evals_per_s": B * (M - 1) * N / (per["fps"] * 1e-3)})
        gf = kern["group_feat"]
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("group_feat_dram_bytes")
        except Exception:
            pass
        line = {"metric": METRIC, "value": total_pts / (ms * 1e-3), "unit": "points/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "clouds_per_gpu": B, "sharding": "by cloud, no data-path collective",
                           "l2_policy": "no flush: one step touches 843 MB (inputs 281 MB + outputs 562 MB) > 126 MB L2",
                           "rank_cpu_binding": numa_cpus,
                           "steps_in_flight": NPIPE, "ms_per_step_one_in_flight": ms_serial},
                "e2e": {"value": total_pts / (ms_e2e * 1e-3), "unit": "points/s", "ms_per_step": ms_e2e,
                        "h2d_bytes_per_step": (xyz_h.numel() + feat_h.numel()) * 4,
                        "d2h_bytes_per_step": sum(t.numel() for t in out_h) * 4},
                "gpu_launches": int(launches),
                "roofline": {"kernel": "group_points (features, C=64): dominant HBM-bound kernel; FPS dominates time but is issue/latency-bound, see kernels",
                             "bound": "hbm", "achieved": gf["hbm_gbs"], "peak": peak, "unit": "GB/s", "frac": gf["hbm_frac"],
                             "traffic": traffic, "peak_source": peak_src},
                "kernels": kern, "clocks": sampler.summary() if sampler else None}
    if world > 1:
        dist.barrier()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import cpu_ref as oracle  # checker/baseline leg only; never on the product path
        oracle.load()
        oracle.set_num_threads(os.cpu_count() or 1)
        xs, fs = xyz_h.numpy(), feat_h.numpy()
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
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--in-flight", type=int, default=3, help="independent steps in flight for `value` (streams + output sets)")
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
