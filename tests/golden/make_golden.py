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
    # the protein models BASELINE configs 3 and 4 name, from the published tables (iqtree2_b200/data/protein_models.json);
    # their eigen-systems come from the REFERENCE's own solver (EigenDecomposition::eigensystem_sym in oracle/_ref)
    "aa_lg_g4": (dict(ntaxa=9, nsite=90, nstates=20, seed=14, gaps=0.03, model="LG"), False, 0.0, 0),
    "aa_lg_c60_g4": (dict(ntaxa=6, nsite=40, nstates=20, nmix=60, seed=15, model="LG+C60"), False, 0.0, 2),
}
NO_PARTIALS = ("dna_deep_scaling", "example_phy", "aa_lg_c60_g4")


def reference_solved_model(name):
    """LG / LG+C60 with every class decomposed by the reference's own eigen-solver."""
    from tests.test_eigen import ref_eigen
    R = W.lg_exchangeabilities()
    exch = R[np.triu_indices(20, 1)]
    d = W.protein_data()
    if name == "LG":
        classes, mixw = [np.asarray(d["LG"]["freqs"])], np.ones(1)
    else:
        classes, mixw = [np.asarray(p) for p in d["C60"]["profiles"]], np.asarray(d["C60"]["weights"])
    ev, U, Ui, F = [], [], [], []
    for fr in classes:
        fr = fr / fr.sum()
        e, u, ui = ref_eigen(exch, fr, 1.0)
        ev.append(e); U.append(u); Ui.append(ui); F.append(fr)
    return W.EigenModel(nstates=20, eval=np.ascontiguousarray(ev), evec=np.ascontiguousarray(U),
                        inv_evec=np.ascontiguousarray(Ui), freqs=np.ascontiguousarray(F), mixw=mixw / mixw.sum(),
                        Q=np.zeros(0))


def problem(name):
    kw = CASES[name][0]
    if kw is None:
        aln = W.alignment_from_phylip(os.path.join(HERE, "example.phy"))
        tree = W.Tree.from_newick(open(os.path.join(HERE, "example17.nwk")).read(), aln.names)
        rates, props = W.discrete_gamma(0.5)
        return tree, W.gtr_model(), rates, props, aln
    kw = dict(kw)
    named = kw.pop("model", None)
    tree, model, rates, props, aln = make_problem(**kw)
    if named:
        model = reference_solved_model(named)
        aln = W.simulate_alignment(tree, model, rates, props, kw["nsite"], kw["seed"] + 7, compress=True)
    return tree, model, rates, props, aln


def main():
    only = sys.argv[1:]  # case names; none = all
    for name, (_, safe, pinv, leaf) in CASES.items():
        if only and name not in only:
            continue
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
            if name not in NO_PARTIALS:
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
