"""Multi-GPU path on CPU: pattern sharding + the scalar all-reduce, world_size 2 over gloo, with the CPU oracle
standing in for the device kernels (tests only).  The sum over shards must equal the single-process result."""
import os
import socket
import sys

import numpy as np
import pytest

from iqtree2_b200 import build_host, multigpu as MG

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module", autouse=True)
def _host():
    build_host.build()


def test_shard_ranges_cover_patterns_in_whole_tiles():
    for nptn in (0, 1, 31, 32, 33, 1000, 1_000_000, 12345):
        for world in (1, 2, 3, 4, 8):
            prev = 0
            sizes = []
            for r in range(world):
                b, e = MG.shard_range(nptn, world, r)
                assert b == prev and e >= b
                assert b % 32 == 0 or b == nptn
                prev = e
                sizes.append(e - b)
            assert prev == nptn
            assert max(sizes) - min(sizes) < 64  # one tile of imbalance plus the ragged last tile


def test_partition_placement_is_lpt():
    rng = np.random.default_rng(0)
    cost = rng.uniform(500, 1500, size=1000) * 100 * rng.choice([4, 20], size=1000)
    for world in (1, 2, 4, 8):
        g = MG.place_partitions(cost, world)
        load = np.bincount(g, weights=cost, minlength=world)
        assert load.sum() == pytest.approx(cost.sum())
        # LPT bound: no GPU exceeds the mean by more than the largest single item
        assert load.max() <= cost.sum() / world + cost.max()
        assert load.max() / load.mean() < 1.02


def _worker(rank, world, port, oracle_lib, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from iqtree2_b200 import workload as W
    from iqtree2_b200.capi import INTERNAL, LEAF
    from tests.common import make_problem
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        tree, model, rates, props, aln = make_problem(12, 700, seed=9, gaps=0.02, compress=False)
        se = MG.ShardedEngine(aln, model, rates, props, nslots=tree.ntaxa, rank=rank, world=world, dist=dist,
                              lib_path=oracle_lib, prefix="orc_")
        c = tree.adj[0][0]
        tr = W.traversal(tree, c, 0)
        res = se.evaluate(tr, (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c))
        plh = se.pattern_lh()
        q.put((rank, res, (se.begin, se.end), plh))
    finally:
        dist.destroy_process_group()


def test_two_rank_pattern_shards_sum_to_the_whole(oracle_lib):
    import multiprocessing as mp
    import oracle
    from iqtree2_b200 import workload as W
    from iqtree2_b200.capi import INTERNAL, LEAF
    from tests.common import make_problem
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, oracle_lib, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    tree, model, rates, props, aln = make_problem(12, 700, seed=9, gaps=0.02, compress=False)
    eng = oracle.oracle_engine(aln, model, rates, props, nslots=tree.ntaxa)
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    want = eng.evaluate(tr, (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c))
    want_plh = eng.fetch_pattern_lh()
    ranges = sorted(g[2] for g in got)
    assert ranges[0][0] == 0 and ranges[0][1] == ranges[1][0] and ranges[1][1] == aln.nptn
    for rank, res, _, plh in got:
        # every rank holds the same reduced scalars
        for a, b in zip(res, want):
            assert abs(a - b) <= 1e-9 * max(1.0, abs(b))
        assert np.allclose(plh, want_plh, rtol=1e-12, atol=0)


@pytest.mark.gpu
def test_nccl_sharded_engine_matches_single_gpu():
    """ShardedEngine's NCCL path on real devices -- the all-reduce must be ordered after the
    engine's kernels (the engine runs on the stream NCCL is enqueued on) and equal the single-GPU result."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    import os
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29571", os.path.join(here, "nccl_shard_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
