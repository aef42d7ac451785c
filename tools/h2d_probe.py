#!/usr/bin/env python
"""Run under torchrun with N ranks (one per GPU): every rank copies the bench's per-step input (281 MB of pinned host
memory) to its GPU at the same time; prints GB/s per rank for (a) one cudaMemcpyAsync per tensor on one stream,
(b) two copy streams with 4 chunks per tensor.  Attributes the end-to-end scaling of bench.py to the host side."""
import json, os, sys
import torch, torch.distributed as dist
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    sys.stdout.flush(); saved = os.dup(1); os.dup2(2, 1)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local)); dist.barrier()
    sys.stdout.flush(); os.dup2(saved, 1); os.close(saved)
nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 281018368
h = torch.empty(nbytes, dtype=torch.uint8).pin_memory(); h.fill_(1)
d = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
def bar():
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
def run(fn, it=10):
    for _ in range(2): fn()
    bar()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); bar()
    t = torch.tensor([e0.elapsed_time(e1) / it], device="cuda")
    if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return nbytes / (float(t) * 1e-3) / 1e9
one = run(lambda: d.copy_(h, non_blocking=True))
s2 = [torch.cuda.Stream(), torch.cuda.Stream()]
def chunked():
    cur = torch.cuda.current_stream()
    for s in s2: s.wait_stream(cur)
    c = nbytes // 8
    for i in range(8):
        with torch.cuda.stream(s2[i & 1]):
            d[i * c:(i + 1) * c].copy_(h[i * c:(i + 1) * c], non_blocking=True)
    for s in s2: cur.wait_stream(s)
two = run(chunked)
if rank == 0:
    print(json.dumps({"ranks": world, "bytes_per_rank": nbytes, "gbs_per_rank_one_stream": one, "gbs_per_rank_two_streams_8_chunks": two,
                      "aggregate_gbs": max(one, two) * world, "host_cpus": os.cpu_count()}), flush=True)
if world > 1: dist.destroy_process_group()
