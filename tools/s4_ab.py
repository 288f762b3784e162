"""A/B of build/launch variants of the 4-state traversal on ONE box, one process, same alignment: the hooks are read
when a context is created, so each variant gets its own engine.

    python tools/s4_ab.py [--workload cfg2] [--steps 20] [--rounds 2] VARIANT [VARIANT ...]

VARIANT is a comma list of NAME=VALUE hooks ("" or "default" = none), e.g.
    default B200LK_S4_BULK=0 B200LK_S4_RING=2,B200LK_S4_MINB=10
Prints, per variant and round, ms per evaluation (CUDA events on the engine's stream), ms inside the partial-update
launches, and the three sums (which must not differ between variants by a single bit).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402  (also restores the CPU affinity libgomp narrows)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--patterns", type=int, default=0)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--rounds", type=int, default=2)
    ap.add_argument("variants", nargs="+")
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
    hooks = [k for k in os.environ if k.startswith("B200LK_")]
    out = []
    for rnd in range(args.rounds):
        for v in args.variants:
            for k in hooks + [k for k in os.environ if k.startswith("B200LK_")]:
                os.environ.pop(k, None)
            if v not in ("", "default"):
                for kv in v.split(","):
                    k, val = kv.split("=")
                    os.environ[k] = val
            eng = engine_for(aln, model, rates, props, nslots=tree.ntaxa, device=0)
            eng.set_stream(stream.cuda_stream)
            for _ in range(3):
                eng.evaluate_async(packed, dad, node, L)
            torch.cuda.synchronize()
            eng.get_profile(reset=True)
            eng.set_profiling(True)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(args.steps):
                eng.evaluate_async(packed, dad, node, L)
            e1.record(stream)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.steps
            eng.set_profiling(False)
            pms, pl = eng.get_profile(reset=True)
            res = eng.read_results()
            row = {"variant": v or "default", "round": rnd, "ms_per_eval": round(ms, 4),
                   "partial_ms_per_eval": round(pms / args.steps, 4), "partial_launches": pl // args.steps,
                   "sums_hex": [float(x).hex() for x in res]}
            print(json.dumps(row), flush=True)
            out.append(row)
            eng.close()
    ref = out[0]["sums_hex"]
    same = all(r["sums_hex"] == ref for r in out)
    print(json.dumps({"all_variants_bit_identical": same}))
    return 0 if same else 1


if __name__ == "__main__":
    sys.exit(main())
