"""CPU tests of the oracle restatement: internal consistency properties that any
correct implementation of the path satisfies."""
import numpy as np
import pytest

import oracle
from iqtree2_b200 import workload as W
from iqtree2_b200.capi import INTERNAL, LEAF
from tests.common import full_eval, make_problem, rel


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("nstates,nmix", [(4, 1), (20, 1), (5, 2)])
def test_rerooting_invariance_and_derivatives(oracle_lib, safe, nstates, nmix):
    tree, model, rates, props, aln = make_problem(9, 300, nstates=nstates, nmix=nmix, seed=3)
    oracle.set_ref_quirks(False)   # odd nstates: see dot_triple in phylokernel_oracle.c
    eng = oracle.oracle_engine(aln, model, rates, props, nslots=16, safe_numeric=safe)
    vals = []
    for leaf in (0, 4, 8):
        r = full_eval(eng, tree, leaf)
        vals.append(r["lnl"])
        assert rel(r["lnl"], r["lnl_buf"]) < 1e-12
        assert rel(r["lnl"], float((r["pattern_lh"] * aln.ptn_freq).sum())) < 1e-12
        # finite differences of lnL(t) on this branch
        L = tree.get_length(leaf, tree.adj[leaf][0])
        h = 1e-5
        lp, lm = eng.lnl_from_buffer(L + h), eng.lnl_from_buffer(L - h)
        assert abs((lp - lm) / (2 * h) - r["df"]) <= 1e-5 * max(1.0, abs(r["df"]))
        assert abs((lp - 2 * r["lnl"] + lm) / h / h - r["ddf"]) <= 2e-3 * max(1.0, abs(r["ddf"]))
    assert max(vals) - min(vals) < 1e-9 * abs(vals[0])


def test_norm_and_safe_agree_and_scaling_fires(oracle_lib):
    # long tree: likelihoods underflow 2^-256 so that scale counts are non-zero
    tree, model, rates, props, aln = make_problem(300, 150, seed=5, mean_len=0.4)
    res = {}
    for safe in (False, True):
        eng = oracle.oracle_engine(aln, model, rates, props, nslots=300, safe_numeric=safe)
        res[safe] = full_eval(eng, tree, 0)
    assert max(int(s.max()) for s in res[False]["scale"].values()) >= 1
    assert rel(res[False]["lnl"], res[True]["lnl"]) < 1e-10
    assert rel(res[False]["df"], res[True]["df"]) < 1e-8


def test_internal_branch_matches_leaf_branch(oracle_lib):
    tree, model, rates, props, aln = make_problem(10, 200, seed=2)
    eng = oracle.oracle_engine(aln, model, rates, props, nslots=40)
    ref = full_eval(eng, tree, 0)["lnl"]
    u = tree.ntaxa
    v = [w for w in tree.adj[u] if not tree.is_leaf(w)][0]
    t1 = W.traversal(tree, u, v)
    t2 = [(a, b, s + 20, [(k, (i + 20 if k == INTERNAL else i), l) for k, i, l in ch])
          for a, b, s, ch in W.traversal(tree, v, u)]
    eng.update_partials(t1)
    eng.update_partials(t2)
    lnl = eng.branch_lnl((INTERNAL, t2[-1][2]), (INTERNAL, t1[-1][2]), tree.get_length(u, v))
    assert rel(lnl, ref) < 1e-12


def test_multifurcating_equals_resolved_with_zero_branch(oracle_lib):
    # a polytomy is the limit of a binary resolution with a zero-length internal branch
    nwk_poly = "(A:0.1,B:0.2,(C:0.1,D:0.3,E:0.2,F:0.1):0.15);"
    nwk_bin = "(A:0.1,B:0.2,(C:0.1,(D:0.3,(E:0.2,F:0.1):0.0):0.0):0.15);"
    names = list("ABCDEF")
    model = W.gtr_model()
    rates, props = W.discrete_gamma(0.7)
    tb = W.Tree.from_newick(nwk_bin, names)
    aln = W.simulate_alignment(tb, model, rates, props, 150, 11, compress=True)
    out = []
    for nwk in (nwk_poly, nwk_bin):
        t = W.Tree.from_newick(nwk, names)
        eng = oracle.oracle_engine(aln, model, rates, props, nslots=8)
        out.append(full_eval(eng, t, 0)["lnl"])
    assert rel(out[0], out[1]) < 1e-12


def test_invariant_sites_and_ambiguity(oracle_lib):
    tree, model, rates, props, aln = make_problem(8, 300, seed=9, gaps=0.1)
    e0 = oracle.oracle_engine(aln, model, rates, props, nslots=8, p_invar=0.0)
    e1 = oracle.oracle_engine(aln, model, rates, props, nslots=8, p_invar=0.2)
    a, b = full_eval(e0, tree, 0), full_eval(e1, tree, 0)
    assert np.isfinite(a["lnl"]) and np.isfinite(b["lnl"]) and a["lnl"] != b["lnl"]


def test_reference_quirk_odd_states_double_counts_last_entry(oracle_lib):
    """Documents tree/phylokernelnew.h:2491 + :700-724: with odd nstates and even ncat_mix the
    reference's dotProductTriple adds the last block entry twice, so its df differs from the
    true derivative.  The CUDA path computes the true derivative (deviations)."""
    tree, model, rates, props, aln = make_problem(7, 200, nstates=5, nmix=1, seed=4)
    out = {}
    for q in (True, False):
        oracle.set_ref_quirks(q)
        eng = oracle.oracle_engine(aln, model, rates, props, nslots=8)
        out[q] = full_eval(eng, tree, 0)
    oracle.set_ref_quirks(True)
    assert out[True]["lnl"] == out[False]["lnl"]
    assert out[True]["df"] != out[False]["df"]
