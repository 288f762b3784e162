"""Run under torchrun (tests/test_multigpu.py::test_nccl_sharded_engine_matches_single_gpu): every rank holds a pattern
shard of the same alignment (iqtree2_b200.multigpu.ShardedEngine), the NCCL all-reduce of [lnL, df, ddf] runs on the
engine's stream, and rank 0 compares the reduced values with a single-GPU evaluation of the whole alignment."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from iqtree2_b200 import multigpu as MG, workload as W  # noqa: E402
from iqtree2_b200.capi import INTERNAL, LEAF, engine_for  # noqa: E402
from tests.common import make_problem  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    tree, model, rates, props, aln = make_problem(40, 200_000, seed=77, compress=False)
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    args = (tr, (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c))
    se = MG.ShardedEngine(aln, model, rates, props, nslots=40, rank=rank, world=world, dist=dist, device=local)
    worst = 0.0
    got = None
    for _ in range(20):  # repeated: a stream race between the kernels and the all-reduce shows as a changing result
        r = se.evaluate(*args)
        if got is not None:
            worst = max(worst, max(abs(a - b) for a, b in zip(r, got)))
        got = r
    # the same sums exchanged inside the finishing kernel over NVLink peer memory (b200lk_peer_connect): no collective call
    b, e = MG.shard_range(aln.nptn, world, rank)
    sub = W.Alignment(aln.nstates, aln.names, np.ascontiguousarray(aln.states[:, b:e]), np.ascontiguousarray(aln.ptn_freq[b:e]))
    pe = engine_for(sub, model, rates, props, nslots=40, device=local)
    pe.connect_peers(dist)
    peer_got, peer_worst = None, 0.0
    for _ in range(20):
        r = pe.evaluate(*args)
        if peer_got is not None:
            peer_worst = max(peer_worst, max(abs(a - c2) for a, c2 in zip(r, peer_got)))
        peer_got = r
    # every rank must hold bit-identical totals (rank-ordered sum)
    mine = torch.tensor(peer_got, dtype=torch.float64, device="cuda")
    lo, hi = mine.clone(), mine.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    same_everywhere = bool(torch.equal(lo, hi))
    ok = True
    if rank == 0:
        one = engine_for(aln, model, rates, props, nslots=40, device=local)
        want = one.evaluate(*args)
        err = max(abs(a - b) / max(1.0, abs(b)) for a, b in zip(got, want))
        print("nccl sharded: got %r want %r rel %.2e repeat-diff %.2e" % (got, want, err, worst), flush=True)
        perr = max(abs(a - b2) / max(1.0, abs(b2)) for a, b2 in zip(peer_got, want))
        print("peer exchange: got %r rel %.2e repeat-diff %.2e identical on all ranks: %s" % (peer_got, perr, peer_worst, same_everywhere),
              flush=True)
        ok = err <= 1e-9 and worst == 0.0 and perr <= 1e-9 and peer_worst == 0.0 and same_everywhere
    flag = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(flag)
    dist.destroy_process_group()
    sys.exit(int(flag.item()) and 1)


if __name__ == "__main__":
    main()
