#!/usr/bin/env python
"""One `optimizeAllBranches` sweep (tree/phylotree.cpp:2689-2765 -> optimizeOneBranch :2585-2635 -> minimizeNewton,
utils/optimization.cpp:422-499) on the cfg2 workload: the caller that issues most of the path's calls in a real run.

GPU: b200::PhyloTree through the host mirror -- per branch one partial update for the re-rooting, theta, and the
device Newton-Raphson loop (one 32-byte read-back per branch).  CPU: the reference's own PhyloTree (oracle/_ref,
AVX+FMA kernel, OpenMP over all host cores) on the first --cpu-patterns patterns of the same alignment.

    python tools/sweep_bench.py [--workload cfg2] [--patterns N] [--cpu-patterns M] [--sweeps K] > line.json
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--patterns", type=int, default=0)
    ap.add_argument("--cpu-patterns", type=int, default=100000)
    ap.add_argument("--sweeps", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import torch
    import bench
    from iqtree2_b200 import build_host, host as H, workload as W
    if not torch.cuda.is_available():
        raise SystemExit("sweep_bench.py needs a CUDA device: there is no CPU fallback on this path")
    build_host.build()
    T, N, n, *_ = bench.WORKLOADS[args.workload]
    N = args.patterns or N
    tree, model, rates, props, aln = bench.build_workload(args.workload, N)
    names = tree.names
    tree.names = ["T%d" % i for i in range(T)]
    newick = tree.to_newick(0, 17)
    tree.names = names
    ha = H.HostAlignment.from_states(aln.states, n, compress=False)
    t = H.HostTree(newick, ha, H.HostModel.from_eigen(model), H.HostRate.categories(rates, props))
    start = t.lengths()
    lnl0 = t.compute_likelihood()
    t.optimize_all_branches(iterations=1)  # warm-up: allocations, first launches
    out = {"workload": args.workload, "ntaxa": T, "patterns": int(aln.nptn), "branches": int(t.nbranches),
           "lnl_start": lnl0, "gpu": []}
    for k in range(1, args.sweeps + 1):
        t.set_lengths(start)
        t.clear_partials()
        t.compute_likelihood()
        torch.cuda.synchronize()
        l0 = t.gpu_launches() if hasattr(t, "gpu_launches") else 0
        t0 = time.perf_counter()
        lnl = t.optimize_all_branches(iterations=1)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        out["gpu"].append({"ms": dt * 1e3, "lnl_after_one_sweep": lnl,
                           "launches": (t.gpu_launches() - l0) if hasattr(t, "gpu_launches") else None})
    best = min(g["ms"] for g in out["gpu"])
    out["gpu_ms_per_sweep"] = best
    out["gpu_us_per_branch"] = best * 1e3 / t.nbranches
    out["gpu_pattern_branches_per_sec"] = t.nbranches * aln.nptn / (best * 1e-3)
    if not args.no_cpu:
        import oracle
        if oracle.ref_available():
            M = min(args.cpu_patterns, aln.nptn)
            sub = W.Alignment(aln.nstates, aln.names, np.ascontiguousarray(aln.states[:, :M]),
                              np.ascontiguousarray(aln.ptn_freq[:M]))
            cores = os.cpu_count() or 1
            ref = oracle.RefRun(tree, sub, model, rates, props, kernel=oracle.LK_AVX_FMA, threads=cores)
            ref.compute_likelihood()
            t0 = time.perf_counter()
            rl = ref.optimize_all_branches(iterations=1)
            dt = time.perf_counter() - t0
            out["cpu"] = {"kind": "reference", "cores": cores, "patterns": M, "ms_per_sweep": dt * 1e3,
                          "lnl_after_one_sweep": rl,
                          "pattern_branches_per_sec": t.nbranches * M / dt}
            out["speedup_per_pattern_branch"] = out["gpu_pattern_branches_per_sec"] / out["cpu"]["pattern_branches_per_sec"]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
