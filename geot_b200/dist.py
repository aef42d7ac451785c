"""Multi-GPU plumbing for the hot path: one process per GPU, clouds sharded by rank, NO data-path collective.

Every op of include/geot_b200.h is independent per cloud, so a rank simply runs the single-GPU path on its
contiguous block of clouds (`shard_clouds`).  The only exchange per training step is the transition-matrix
statistics and the T-predictor gradients — a few hundred KB, latency-bound — done with
torch.distributed (NCCL over NVLink on the GPU box, gloo in the CPU tests):

  combine_class_T      per-class anchor rows: all_gather (value | row), keep the arg-max (ties: lowest rank —
                       identical to the reference's single-process first-max over the concatenated batch,
                       train.py:509-526, because ranks hold contiguous blocks of clouds in rank order)
  allreduce_class_means  cm / c_num running class means (train.py:868-897): count-weighted combine
  allreduce_grads      one packed fp32 all_reduce(SUM) for a list of parameters' .grad (T predictor: C*C*2C floats),
                       optionally on a side stream so it overlaps the rest of the backward

  StatsExchange        the three above fused around ONE collective per step: a packed record per rank
                       (csrc/stats_exchange.cu: geot_stats_pack), one all_gather_into_tensor on a side stream, one
                       combine kernel (geot_stats_combine: arg-max pick + fixed-rank-order sums) - what a training
                       step uses; the three separate functions stay as the readable definition and for the gloo tests

The reference reduces none of this (single process, train.py:7); these semantics are this project's definition.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def _world(group=None):
    return dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1


def _rank(group=None):
    return dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0


def shard_clouds(num_clouds: int, rank: int | None = None, world: int | None = None):
    """Contiguous block of clouds owned by `rank`: (first, count).  Remainders go to the lowest ranks."""
    rank = _rank() if rank is None else rank
    world = _world() if world is None else world
    base, rem = divmod(num_clouds, world)
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


def shard_offsets(offset, rank: int | None = None, world: int | None = None):
    """Offset-format twin: cumulative `offset` (b,) of the whole batch -> (first_cloud, count, point_lo, point_hi,
    local cumulative offset rebased to this rank's first point).  Global indices returned by the offset ops on the
    shard are local-to-shard; add point_lo to make them global again."""
    o = offset.detach().cpu().to(torch.int64)
    first, cnt = shard_clouds(o.numel(), rank, world)
    lo = int(o[first - 1]) if first > 0 else 0
    hi = int(o[first + cnt - 1]) if cnt > 0 else lo
    local = (o[first:first + cnt] - lo).to(torch.int32).to(offset.device)
    return first, cnt, lo, hi, local


@torch.no_grad()
def combine_class_T(class_T, best_val, group=None):
    """class_T (C,C) rows picked on this rank's shard with their scores best_val (C) -> the global pick."""
    W = _world(group)
    if W == 1:
        return class_T
    C = class_T.shape[0]
    packed = torch.cat([best_val.view(C, 1), class_T], dim=1).contiguous()  # (C, C+1)
    gathered = [torch.empty_like(packed) for _ in range(W)]
    dist.all_gather(gathered, packed, group=group)
    stack = torch.stack(gathered)                      # (W, C, C+1)
    winner = stack[:, :, 0].argmax(dim=0)              # first maximum = lowest rank on ties
    return stack[winner, torch.arange(C, device=stack.device), 1:].contiguous()


@torch.no_grad()
def allreduce_class_means(cm, c_num, group=None):
    """Running class means cm (C,C) with counts c_num (C) -> count-weighted global means, summed counts (in place)."""
    if _world(group) == 1:
        return cm, c_num
    C = cm.shape[0]
    buf = torch.cat([(cm * c_num.view(C, 1)).reshape(-1), c_num.reshape(-1)]).contiguous()
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    tot = buf[C * C:]
    cm.copy_(torch.where(tot.view(C, 1) > 0, buf[:C * C].view(C, C) / tot.clamp_min(1e-30).view(C, 1), cm))
    c_num.copy_(tot)
    return cm, c_num


class GradAllReduce:
    """Packed SUM all-reduce of parameter gradients (T predictor / T_linear / sigma) on a side stream.

        h = GradAllReduce(params); loss.backward(); h.launch(); ...; h.wait(); optimizer.step()
    """

    def __init__(self, params, group=None, average=True):
        self.params = [p for p in params if p.requires_grad]
        self.group, self.average = group, average
        self._buf = None
        self._stream = None
        self._work = None

    def launch(self):
        if _world(self.group) == 1 or not self.params:
            return self
        grads = [(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in self.params]
        dev = grads[0].device
        if dev.type == "cuda":
            if self._stream is None:
                self._stream = torch.cuda.Stream(device=dev)
            self._stream.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(self._stream):
                self._buf = torch.cat(grads)
                dist.all_reduce(self._buf, op=dist.ReduceOp.SUM, group=self.group)
        else:
            self._buf = torch.cat(grads)
            dist.all_reduce(self._buf, op=dist.ReduceOp.SUM, group=self.group)
        return self

    def wait(self):
        if self._buf is None:
            return
        if self._stream is not None:
            torch.cuda.current_stream(self._buf.device).wait_stream(self._stream)
        scale = 1.0 / _world(self.group) if self.average else 1.0
        off = 0
        for p in self.params:
            n = p.numel()
            g = self._buf[off:off + n].view_as(p) * scale
            if p.grad is None:
                p.grad = g.clone()
            else:
                p.grad.copy_(g)
            off += n
        self._buf = None


class StatsExchange:
    """All per-step statistics of the transition-matrix path in ONE collective (CUDA tensors; NCCL).

        ex = StatsExchange(C, params=t_predictor.parameters())
        ...
        ex.launch(class_T_local, best_val_local, cm, c_num)     # after backward: packs (incl. the params' .grad),
                                                                 # all-gathers on a side stream
        ...                                                      # anything that does not need the results
        class_T, cm, c_num = ex.wait()                           # combined values; params' .grad now hold the mean

    Equivalent to combine_class_T + allreduce_class_means + GradAllReduce (tests/test_multidevice_gpu.py), with
    6 launches and one collective per step instead of ~50 tensor ops and three.  World size 1: a pass-through."""

    def __init__(self, num_classes, params=None, group=None, average=True):
        self.C = int(num_classes)
        self.params = [p for p in (params or []) if p.requires_grad]
        self.P = sum(p.numel() for p in self.params)
        self.group, self.average = group, average
        self._stream = None
        self._rec = self._all = self._out = None
        self._pending = None

    def _buffers(self, dev):
        from . import _lib
        if self._rec is None or self._rec.device != dev:
            W = _world(self.group)
            L = int(_lib.lib().geot_stats_record_floats(self.C, self.P))
            self._rec = torch.empty((L,), dtype=torch.float32, device=dev)
            self._all = torch.empty((W, L), dtype=torch.float32, device=dev)
            self._grads = torch.empty((max(self.P, 1),), dtype=torch.float32, device=dev)
            self._stream = torch.cuda.Stream(device=dev)
        return self._rec, self._all

    def launch(self, class_T, best_val, cm, c_num):
        from . import _lib
        W = _world(self.group)
        if W == 1:
            self._pending = (class_T, cm, c_num, None)
            return self
        dev = _lib.check_cuda(class_T, best_val, cm, c_num)
        rec, gathered = self._buffers(dev)
        cur = torch.cuda.current_stream(dev)
        self._stream.wait_stream(cur)
        with _lib.on_device(dev), torch.cuda.stream(self._stream):
            if self.P:
                views = [(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in self.params]
                torch.cat(views, out=self._grads)
            _lib.call("geot_stats_pack", self.C, self.P, _lib.ptr(class_T), _lib.ptr(best_val), _lib.ptr(cm), _lib.ptr(c_num),
                      _lib.ptr(self._grads) if self.P else None, _lib.ptr(rec), _lib.stream_ptr())
            dist.all_gather_into_tensor(gathered.view(-1), rec, group=self.group)
            out_T, out_cm, out_cn = torch.empty_like(class_T), cm.clone(), torch.empty_like(c_num)
            _lib.call("geot_stats_combine", W, self.C, self.P, _lib.ptr(gathered), 1.0 / W if self.average else 1.0,
                      _lib.ptr(out_T), _lib.ptr(out_cm), _lib.ptr(out_cn), _lib.ptr(self._grads) if self.P else None,
                      _lib.stream_ptr())
        for t in (class_T, best_val, cm, c_num):
            t.record_stream(self._stream)
        self._pending = (out_T, out_cm, out_cn, dev)
        return self

    def wait(self):
        assert self._pending is not None, "StatsExchange.wait() without launch()"
        out_T, out_cm, out_cn, dev = self._pending
        self._pending = None
        if dev is None:
            return out_T, out_cm, out_cn
        torch.cuda.current_stream(dev).wait_stream(self._stream)
        if self.P:
            off, views = 0, []
            for p in self.params:
                views.append(self._grads[off:off + p.numel()].view_as(p))
                off += p.numel()
            missing = [p for p in self.params if p.grad is None]
            for p in missing:
                p.grad = torch.empty_like(p)
            torch._foreach_copy_([p.grad for p in self.params], views)
        return out_T, out_cm, out_cn
