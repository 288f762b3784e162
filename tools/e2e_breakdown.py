"""Where the host time of one blocking evaluation goes (the e2e figure of bench.py): set_model, the asynchronous
enqueue (node program + launches), and the wait for the three sums -- next to the device-only time of the same step.

    python tools/e2e_breakdown.py [--workload cfg2] [--steps 30]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--patterns", type=int, default=0)
    ap.add_argument("--steps", type=int, default=30)
    args = ap.parse_args()
    import torch
    from iqtree2_b200 import workload as W
    from iqtree2_b200.capi import INTERNAL, LEAF, engine_for, pack_traversal
    tree, model, rates, props, aln, _ = bench.build_workload(args.workload, args.patterns or None)
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    packed = pack_traversal(tr)
    dad, node, L = (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    eng = engine_for(aln, model, rates, props, nslots=tree.ntaxa, device=0)
    eng.set_stream(stream.cuda_stream)
    import numpy as np
    margs = (model.eval, model.evec, model.inv_evec, np.asarray(rates), np.asarray(props), model.mixw, W.tip_partial_lh(model))
    for _ in range(5):
        eng.evaluate_async(packed, dad, node, L)
        eng.read_results()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        eng.evaluate_async(packed, dad, node, L)
    e1.record(stream)
    torch.cuda.synchronize()
    dev_ms = e0.elapsed_time(e1) / args.steps
    t_model = t_enq = t_wait = 0.0
    for _ in range(args.steps):
        a = time.perf_counter()
        if margs is not None:
            eng.set_model(*margs)
        b = time.perf_counter()
        eng.evaluate_async(packed, dad, node, L)
        cc = time.perf_counter()
        eng.read_results()
        d = time.perf_counter()
        t_model += b - a
        t_enq += cc - b
        t_wait += d - cc
    k = 1e3 / args.steps
    print(json.dumps({"device_ms_per_eval": round(dev_ms, 4), "set_model_ms": round(t_model * k, 4),
                      "enqueue_ms": round(t_enq * k, 4), "wait_ms": round(t_wait * k, 4),
                      "e2e_ms": round((t_model + t_enq + t_wait) * k, 4), "set_model_timed": margs is not None}))


if __name__ == "__main__":
    main()
