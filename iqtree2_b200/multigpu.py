"""Pattern sharding and partition placement over the GPUs of one box, one process per GPU.

The path shards by site pattern: rank r owns a contiguous range of patterns (whole 32-pattern
tiles, b200::shardPatterns) of EVERY partial vector; nothing is exchanged during a traversal; the only collective
is the sum of [lnL, df, ddf] (3 doubles) -- on the device result buffer over NCCL on B200, or on a host tensor over
gloo in the CPU tests.  Supermatrices place whole partitions by cost instead (b200::placePartitions).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import host as H


def shard_range(nptn, world, rank, tile=32):
    b, e = C.c_int64(), C.c_int64()
    if H.lib().b200host_shard_patterns(C.c_int64(nptn), world, rank, tile, C.byref(b), C.byref(e)) != 0:
        raise ValueError("bad shard arguments")
    return b.value, e.value


def place_partitions(costs, world):
    costs = np.ascontiguousarray(costs, dtype=np.float64)
    out = np.empty(len(costs), dtype=np.int32)
    if H.lib().b200host_place_partitions(costs.ctypes.data_as(C.c_void_p), len(costs), world,
                                         out.ctypes.data_as(C.c_void_p)) != 0:
        raise ValueError("bad placement arguments")
    return out


class ShardedEngine:
    """One rank's share of a likelihood problem: an Engine (capi) over this rank's pattern range, and the
    all-reduce that turns per-shard sums into the tree's lnL / df / ddf."""

    def __init__(self, aln, model, rates, props, *, nslots, rank, world, dist=None, device=None,
                 lib_path=None, prefix="b200lk_", safe_numeric=False, p_invar=0.0):
        from . import workload as W
        from .capi import LIB_PATH, engine_for
        self.rank, self.world, self.dist = rank, world, dist
        self.total_patterns = aln.nptn
        self.begin, self.end = shard_range(aln.nptn, world, rank)
        sub = W.Alignment(aln.nstates, aln.names, np.ascontiguousarray(aln.states[:, self.begin:self.end]),
                          np.ascontiguousarray(aln.ptn_freq[self.begin:self.end]))
        self.local = sub
        self.eng = engine_for(sub, model, rates, props, nslots=nslots, p_invar=p_invar, safe_numeric=safe_numeric,
                              device=0 if device is None else device, lib_path=lib_path or LIB_PATH, prefix=prefix)
        self.on_device = prefix == "b200lk_"
        self._res = self.eng.result_tensor() if self.on_device else None
        self._stream = None
        if self.on_device and world > 1:
            # The context's own stream is cudaStreamNonBlocking: a collective enqueued on another stream is NOT ordered
            # after the lnL / derv kernels.  The engine and NCCL share ONE dedicated stream (not torch's default stream:
            # its handle is 0, which b200lk_set_stream reads as "restore the context's own stream").
            import torch
            self._stream = torch.cuda.Stream(device=self.eng.dims.device)
            self.eng.set_stream(self._stream.cuda_stream)

    def evaluate(self, entries, dad, node, length):
        """Full evaluation on this shard, then the sum over shards.  Returns (lnL, df, ddf) of the whole alignment."""
        if self.on_device:
            from .capi import pack_traversal
            self.eng.evaluate_async(pack_traversal(entries), dad, node, length)
            if self.world > 1:
                import torch
                with torch.cuda.stream(self._stream):
                    self.dist.all_reduce(self._res)  # NCCL, in place on the device result buffer, same stream
                    out = self._res.to("cpu", non_blocking=False)
                return tuple(float(x) for x in out)
            return self.eng.read_results()
        import torch
        part = torch.tensor(self.eng.evaluate(entries, dad, node, length), dtype=torch.float64)
        if self.world > 1:
            self.dist.all_reduce(part)
        return tuple(float(x) for x in part)

    def pattern_lh(self):
        """Per-pattern lnL of the whole alignment, gathered in pattern order (what -wsl / UFBoot read)."""
        mine = self.eng.fetch_pattern_lh()
        if self.world == 1:
            return mine
        import torch
        sizes = [e - b for b, e in (shard_range(self.total_patterns, self.world, r) for r in range(self.world))]
        pad = max(sizes)
        send = torch.zeros(pad, dtype=torch.float64)
        send[:len(mine)] = torch.from_numpy(mine)
        bufs = [torch.empty(pad, dtype=torch.float64) for _ in range(self.world)]
        self.dist.all_gather(bufs, send)
        return np.concatenate([bufs[r][:sizes[r]].numpy() for r in range(self.world)])
