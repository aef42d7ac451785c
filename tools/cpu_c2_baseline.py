#!/usr/bin/env python
"""the reference's OWN device-agnostic torch code for the sample+group path, timed on CPU.

Runs only where /root/reference exists (the build container - the GPU box has no copy of the reference), so the number is
"reported only": it is measured on THIS machine's cores, not on the GPU box's host, and it is not a parity oracle (the
torch code's semantics differ from the CUDA kernels': random FPS start, `>` instead of `<` at the radius, matmul-expanded
distances, unsorted topk).  The functions are taken from the reference's source files as they are (ast, no package
import: openpoints needs its compiled extensions); the only edit is the one BASELINE.md names - the random FPS start
index (pointmlp.py:98) fixed to 0.

  python tools/cpu_c2_baseline.py [--clouds 2] [--out profiles/r02_cpu_c2_container.json]
"""
import argparse, ast, json, os, sys, time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

REF = os.environ.get("GEOT_REFERENCE", "/root/reference")
N, M, NS, C, RADIUS = 16384, 1024, 32, 64, 0.1


def functions_of(path, names):
    src = open(os.path.join(REF, path)).read()
    out = {}
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            out[node.name] = ast.get_source_segment(src, node)
    missing = set(names) - set(out)
    assert not missing, f"{path}: {missing} not found"
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--clouds", type=int, default=2, help="clouds in the timed sample (the full batch has 64)")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    if not os.path.isdir(REF):
        print(json.dumps({"unavailable": f"{REF} does not exist on this machine"}))
        return
    from geot_b200 import synth
    torch.set_num_threads(os.cpu_count())
    ns = {"torch": torch}
    fn = functions_of("openpoints/models/backbone/pointmlp.py", ["square_distance", "index_points", "farthest_point_sample", "query_ball_point"])
    fps_src = fn["farthest_point_sample"]
    fixed = fps_src.replace("farthest = torch.randint(0, N, (B,), dtype=torch.long).to(device)", "farthest = torch.zeros(B, dtype=torch.long).to(device)")
    assert fixed != fps_src, "pointmlp.py:98 (random start) not found"
    fn["farthest_point_sample"] = fixed
    for name, code in fn.items():
        exec(compile(code, f"<reference pointmlp.py:{name}>", "exec"), ns)
    g = functions_of("openpoints/models/layers/group.py", ["torch_grouping_operation"])
    exec(compile(g["torch_grouping_operation"], "<reference group.py:torch_grouping_operation>", "exec"), ns)

    b = args.clouds
    xyz = torch.from_numpy(synth.tooth_batch(b, N))
    feat = torch.from_numpy(np.random.default_rng(0).standard_normal((b, C, N)).astype(np.float32))
    t = {}

    def timed(name, f):
        t0 = time.perf_counter()
        r = f()
        t[name] = time.perf_counter() - t0
        return r

    with torch.no_grad():
        idx = timed("fps", lambda: ns["farthest_point_sample"](xyz, M))
        centres = timed("gather_centres", lambda: ns["index_points"](xyz, idx))
        nbr = timed("ball_query", lambda: ns["query_ball_point"](RADIUS, NS, xyz, centres))
        xyz_t = xyz.transpose(1, 2).contiguous()
        timed("group_xyz", lambda: ns["torch_grouping_operation"](xyz_t, nbr))
        timed("group_feat", lambda: ns["torch_grouping_operation"](feat, nbr))
    total = sum(t.values())
    row = {"what": "BASELINE.md 2.2 C2: the reference's device-agnostic torch code for the bench step, on CPU",
           "where": "build container (no GPU box copy of the reference exists); reported only",
           "cores": os.cpu_count(), "torch_threads": torch.get_num_threads(), "clouds_timed": b, "points_per_cloud": N,
           "seconds": {k: round(v, 4) for k, v in t.items()}, "seconds_total": round(total, 4),
           "points_per_s": b * N / total, "seconds_per_64_clouds_extrapolated": total * 64 / b,
           "sample": f"{b} of the 64 clouds of the bench batch (same generator), one pass"}
    line = json.dumps(row)
    print(line)
    if args.out:
        open(args.out, "w").write(line + "\n")


if __name__ == "__main__":
    main()
