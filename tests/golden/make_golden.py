#!/usr/bin/env python
"""Generate tests/golden/ref_*.npz from the REFERENCE's own kernels (oracle/_ref, built from the unmodified
sources under /root/reference by oracle/Makefile.ref).  Run in the authoring container only:

    python tests/golden/make_golden.py

Each file holds the complete inputs of one small problem (tip states, pattern weights, tree, eigen-system,
rates) and what the reference's AVX+FMA kernel (-lk FMA) computed for it: total lnL at the branch of leaf
`leaf`, per-pattern lnL, per-category pattern likelihoods, df/ddf, lnL-from-buffer at other branch lengths,
and partial_lh + scale_num of every internal node of the traversal (keyed by the sorted leaf set of the
subtree).  tests/test_golden.py checks the oracle (CPU) and the CUDA path (GPU) against them."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle  # noqa: E402
from iqtree2_b200 import workload as W  # noqa: E402
from tests.common import make_problem  # noqa: E402
from tests.test_oracle_vs_ref import leaf_sets  # noqa: E402

CASES = {
    # name: (kwargs of make_problem, safe, p_invar, leaf)
    "dna_gtr_g4": (dict(ntaxa=12, nsite=400, seed=3, gaps=0.05), False, 0.0, 0),
    "dna_gtr_g4_safe": (dict(ntaxa=12, nsite=400, seed=3, gaps=0.05), True, 0.0, 5),
    "dna_deep_scaling": (dict(ntaxa=150, nsite=60, seed=5, mean_len=0.6), False, 0.0, 0),
    "dna_invar": (dict(ntaxa=10, nsite=300, seed=8), False, 0.25, 2),
    "aa_g4": (dict(ntaxa=8, nsite=80, nstates=20, seed=4, gaps=0.03), False, 0.0, 0),
    "aa_mix3_g4": (dict(ntaxa=6, nsite=50, nstates=20, nmix=3, seed=5), False, 0.0, 1),
    "example_phy": (None, False, 0.0, 0),
}


def problem(name):
    kw = CASES[name][0]
    if kw is None:
        aln = W.alignment_from_phylip(os.path.join(HERE, "example.phy"))
        tree = W.Tree.from_newick(open(os.path.join(HERE, "example17.nwk")).read(), aln.names)
        rates, props = W.discrete_gamma(0.5)
        return tree, W.gtr_model(), rates, props, aln
    return make_problem(**kw)


def main():
    for name, (_, safe, pinv, leaf) in CASES.items():
        tree, model, rates, props, aln = problem(name)
        ref = oracle.RefRun(tree, aln, model, rates, props, kernel=oracle.LK_AVX_FMA, safe=safe, p_invar=pinv)
        out = dict(nstates=aln.nstates, states=aln.states, ptn_freq=aln.ptn_freq, newick=tree.to_newick(0, 17),
                   names=np.array(tree.names), eval=model.eval, evec=model.evec, inv_evec=model.inv_evec,
                   freqs=model.freqs, mixw=model.mixw, rates=rates, props=props, p_invar=pinv, safe=safe,
                   leaf=leaf)
        out["lnl"] = ref.lnl_at_leaf(leaf)
        out["pattern_lh"] = ref.pattern_lh()
        out["pattern_lh_cat"] = ref.pattern_lh_cat()
        c = tree.adj[leaf][0]
        tr = W.traversal(tree, c, leaf)
        for slot, ls in leaf_sets(tree, tr).items():
            part, sc = ref.fetch_branch(ls, leaf)
            out["scale_%d" % slot] = sc
            if name not in ("dna_deep_scaling", "example_phy"):
                out["partial_%d" % slot] = part
        out["df"], out["ddf"] = ref.derv()
        out["lnl_buf"] = ref.from_buffer()
        L = tree.get_length(leaf, c)
        ts = np.array([0.25 * L, 2.0 * L, 1e-6, 4.0])
        dd, ll = [], []
        for t in ts:
            ref.set_current_length(float(t))
            dd.append(ref.derv(again=True))
            ll.append(ref.from_buffer())
        out["other_t"], out["other_derv"], out["other_lnl"] = ts, np.array(dd), np.array(ll)
        ref.close()
        if name in ("example_phy", "dna_gtr_g4", "aa_g4"):
            # branch-length optimisation by the reference's own PhyloTree::optimizeAllBranches (Newton-Raphson,
            # tree/phylotree.cpp:2689-2765) from the stored starting lengths: final lnL, the reference's own root
            # choice lnL (computeLikelihood) and every optimised length keyed by the leaf set behind the branch
            ref = oracle.RefRun(tree, aln, model, rates, props, kernel=oracle.LK_AVX_FMA, safe=safe, p_invar=pinv)
            out["lnl_compute_likelihood"] = ref.compute_likelihood()
            out["opt_lnl"] = ref.optimize_all_branches()
            splits, lens = [], []
            for slot, ls in leaf_sets(tree, W.traversal(tree, tree.adj[0][0], 0)).items():
                splits.append(",".join(str(x) for x in sorted(ls)))
                lens.append(ref.branch_length(ls, 0))
            for t in range(tree.ntaxa):
                splits.append(str(t))
                lens.append(ref.branch_length([t], t) if t != 0 else ref.branch_length([0], 0))
            out["opt_splits"], out["opt_lengths"] = np.array(splits), np.array(lens)
            ref.close()
        path = os.path.join(HERE, "ref_%s.npz" % name)
        np.savez_compressed(path, **out)
        print("%-18s lnL %.10f  -> %s (%d kB)" % (name, out["lnl"], os.path.basename(path), os.path.getsize(path) // 1024))


if __name__ == "__main__":
    main()
