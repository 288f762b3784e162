"""Model-probe throughput: K candidate reversible models on one alignment and tree, as
ModelFinder evaluates them (main/phylotesting.cpp:1563-1680): fresh eigen-system + one full lnL evaluation each.
  a) per candidate: host eigen-solve + b200lk_set_model (n^2 upload) + b200lk_evaluate
  b) per candidate b200lk_set_model_rates (device Jacobi), then ONE batched evaluation over all candidates
  c) the reference's AVX+FMA kernel on the host cores (oracle/_ref), eigen-system given
Prints one JSON line.  Usage: python tools/probe_bench.py [--taxa 100] [--patterns 5000] [--models 24] [--states 4|20]"""
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
    ap.add_argument("--patterns", type=int, default=5000)
    ap.add_argument("--models", type=int, default=24)
    ap.add_argument("--states", type=int, default=4)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    from iqtree2_b200 import workload as W
    from iqtree2_b200.capi import INTERNAL, LEAF, Batch, engine_for
    T, N, K, n = args.taxa, args.patterns, args.models, args.states
    rng = np.random.default_rng(8)
    tree = W.random_tree(T, 8)
    gen = W.gtr_model() if n == 4 else W.random_reversible_model(n, 8)
    rates, props = W.discrete_gamma(0.6)
    aln = W.simulate_alignment(tree, gen, rates, props, N, 9, compress=False)
    tipvec = W.state_indicators(n)
    ne = n * (n - 1) // 2
    cands = [(rng.gamma(2.0, 1.0, ne), rng.dirichlet(np.full(n, 8.0))) for _ in range(K)]
    leaf = 0
    c = tree.adj[leaf][0]
    tr = W.traversal(tree, c, leaf)
    top = tr[-1][2]
    L = tree.get_length(leaf, c)
    child_len = np.tile(np.array([ch[2] for e in tr for ch in e[3]]), (K, 1))
    blen = np.full(K, L)
    engines = [engine_for(aln, gen, rates, props, nslots=T) for _ in range(K)]
    batch = Batch(engines)

    def host_model(ex, fr):
        R = np.zeros((n, n))
        R[np.triu_indices(n, 1)] = ex
        return W.make_model([(R + R.T, fr)])

    def path_a():
        out = []
        for e, (ex, fr) in zip(engines, cands):
            m = host_model(ex, fr)
            e.set_model(m.eval, m.evec, m.inv_evec, rates, props, m.mixw, W.tip_partial_lh(m))
            out.append(e.evaluate(tr, (LEAF, leaf), (INTERNAL, top), L, want_derv=False)[0])
        return np.array(out)

    def path_b():
        for e, (ex, fr) in zip(engines, cands):
            e.set_model_rates(ex[None], fr[None], rates, props, np.ones(1), tipvec)
        return batch.evaluate(tr, child_len, (LEAF, leaf), (INTERNAL, top), blen, want_derv=False)[0]

    res = {}
    for name, fn in (("per_candidate_host_eigen", path_a), ("device_eigen_batched", path_b)):
        fn()
        t0 = time.perf_counter()
        for _ in range(args.reps):
            lnl = fn()
        res[name] = {"ms_per_probe": (time.perf_counter() - t0) * 1e3 / (args.reps * K), "lnl0": float(lnl[0])}
    assert abs(res["per_candidate_host_eigen"]["lnl0"] - res["device_eigen_batched"]["lnl0"]) <= 1e-9 * abs(res["device_eigen_batched"]["lnl0"])
    try:
        import oracle
        if oracle.ref_available():
            m = host_model(*cands[0])
            ref = oracle.RefRun(tree, aln, m, rates, props, kernel=oracle.LK_AVX_FMA, threads=os.cpu_count() or 1)
            ref.lnl_at_leaf(leaf)
            t0 = time.perf_counter()
            for _ in range(3):
                ref.clear_partials() if hasattr(ref, "clear_partials") else None
                ref.lnl_at_leaf(leaf)
            res["reference_fma_kernel"] = {"ms_per_probe": (time.perf_counter() - t0) * 1e3 / 3, "cores": os.cpu_count(),
                                           "note": "eigen-system given; evaluation only"}
            ref.close()
    except Exception as e:  # the reference arm is optional here
        res["reference_fma_kernel"] = {"unavailable": str(e)[:100]}
    print(json.dumps({"workload": "%d candidate %d-state models, %d taxa x %d patterns, +G4" % (K, n, T, N), **res}))


if __name__ == "__main__":
    main()
