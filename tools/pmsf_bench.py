"""Site-specific models (PMSF shape): one exchangeability matrix, a frequency profile per pattern, +G4.
One evaluation = all partials stale -> lnL at a leaf branch -> theta -> df/ddf, through b200lk_evaluate_async.
Reports GPU time per evaluation (CUDA events), algorithmic HBM traffic, and the reference's own SITE_MODEL
AVX+FMA kernel on the host cores on a bounded sample (oracle/_ref).  Prints one JSON line.
Usage: python tools/pmsf_bench.py [--taxa 100] [--patterns 100000] [--states 20] [--cpu-sample 5000]"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--taxa", type=int, default=100)
    ap.add_argument("--patterns", type=int, default=100_000)
    ap.add_argument("--states", type=int, default=20)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--cpu-sample", type=int, default=5000)
    args = ap.parse_args()
    import torch
    from iqtree2_b200 import workload as W
    from iqtree2_b200.capi import INTERNAL, LEAF, pack_traversal
    from tests.common import site_engine
    T, N, n = args.taxa, args.patterns, args.states
    rng = np.random.default_rng(12)
    tree = W.random_tree(T, 12)
    gen = W.random_reversible_model(n, 12)
    rates, props = W.discrete_gamma(0.7)
    aln = W.simulate_alignment(tree, gen, rates, props, N, 13, compress=False)
    # per-pattern models, batched eigen-solve
    R = np.zeros((n, n))
    R[np.triu_indices(n, 1)] = rng.gamma(2.0, 1.0, n * (n - 1) // 2)
    R = R + R.T
    pi = rng.dirichlet(np.full(n, 3.0), size=N)
    Q = R[None] * pi[:, None, :]
    idx = np.arange(n)
    Q[:, idx, idx] = 0.0
    Q[:, idx, idx] = -Q.sum(axis=2)
    Q /= -(pi * Q[:, idx, idx]).sum(axis=1)[:, None, None]
    s = np.sqrt(pi)
    S = Q * s[:, :, None] / s[:, None, :]
    S = 0.5 * (S + S.transpose(0, 2, 1))
    w, V = np.linalg.eigh(S)
    models = (np.ascontiguousarray(w), np.ascontiguousarray(V / s[:, :, None]), np.ascontiguousarray(V.transpose(0, 2, 1) * s[:, None, :]))
    eng = site_engine(aln, models, rates, props, nslots=T)
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    packed = pack_traversal(tr)
    dad, node, L = (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c)
    stream = torch.cuda.Stream()
    eng.set_stream(stream.cuda_stream)
    for _ in range(3):
        eng.evaluate_async(packed, dad, node, L)
    res = eng.read_results()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(args.steps):
        eng.evaluate_async(packed, dad, node, L)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    ncat = len(rates)
    block = n * ncat
    # per pattern-node: the pattern's eigen-system (2 n^2 + n doubles) + children (k_int blocks or n-vectors) + output
    alg = 0
    for e in tr:
        k = sum(1 for ch in e[3] if ch[0] == 1)
        alg += 8 * (2 * n * n + n) + 8 * block * (1 + k) + 8 * n * (2 - k) + (1 + k)
    alg *= N
    out = {"workload": "site-specific models (PMSF shape): %d taxa x %d patterns, %d states, +G4" % (T, N, n),
           "ms_per_eval": ms, "updates_per_sec": len(tr) * N / (ms * 1e-3), "lnl_df_ddf": list(res),
           "algorithmic_gb_per_eval": alg / 1e9, "achieved_gbs": alg / (ms * 1e-3) / 1e9}
    try:
        import oracle
        if oracle.ref_available():
            m = min(args.cpu_sample, N)
            sub = W.Alignment(aln.nstates, aln.names, np.ascontiguousarray(aln.states[:, :m]), np.ascontiguousarray(aln.ptn_freq[:m]))
            ref = oracle.RefRun(tree, sub, gen, rates, props, kernel=oracle.LK_AVX_FMA, threads=os.cpu_count() or 1,
                                site_models=tuple(a[:m] for a in models))
            ref.lnl_at_leaf(0)
            t0 = time.perf_counter()
            for _ in range(3):
                ref.lnl_at_leaf(0)
                ref.derv()
            dt = (time.perf_counter() - t0) / 3
            out["reference_fma_kernel"] = {"updates_per_sec": len(tr) * m / dt, "cores": os.cpu_count(), "sample_patterns": m}
            ref.close()
    except Exception as e:
        out["reference_fma_kernel"] = {"unavailable": str(e)[:120]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
