"""Pins the CPU oracle (oracle/phylokernel_oracle.c) against the REFERENCE's own kernels: the unmodified
translation units of /root/reference compiled into oracle/_ref/libiqtree_ref.so (oracle/Makefile.ref,
ref_harness.cpp).  Skipped where that library has not been built; the committed golden vectors
(tests/golden/*.json, test_golden.py) carry the same pin to machines without /root/reference."""
import numpy as np
import pytest

import oracle
from iqtree2_b200 import workload as W
from iqtree2_b200.capi import INTERNAL, LEAF
from tests.common import full_eval, make_problem, rel

pytestmark = pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref not built")

KERNELS = [oracle.LK_SSE2, oracle.LK_AVX, oracle.LK_AVX_FMA]


def leaf_sets(tree, entries):
    """sorted leaf ids of the subtree at `node` seen from `dad`, for every traversal entry"""
    out = {}

    def rec(v, u):
        if tree.is_leaf(v):
            return [v]
        r = []
        for w in tree.adj[v]:
            if w != u:
                r += rec(w, v)
        return r

    for node, dad, slot, _ in entries:
        out[slot] = sorted(rec(node, dad))
    return out


def compare(tree, aln, model, rates, props, kernel, safe, leaf=0, p_invar=0.0, tol=1e-11, check_nodes=True, fused=False):
    ref = oracle.RefRun(tree, aln, model, rates, props, kernel=kernel, safe=safe, p_invar=p_invar, fused=fused)
    safe_eff = ref.safe_numeric       # forced for nstates not in {4, 20} (tree/phylotreesse.cpp:95-96)
    oracle.set_ref_quirks(True)
    eng = oracle.oracle_engine(aln, model, rates, props, nslots=tree.ntaxa, safe_numeric=safe_eff, p_invar=p_invar,
                               fused=fused)
    lnl = ref.lnl_at_leaf(leaf)
    o = full_eval(eng, tree, leaf)
    assert rel(lnl, o["lnl"]) <= tol
    assert rel(ref.pattern_lh(), o["pattern_lh"]) <= tol
    cat_r, cat_o = ref.pattern_lh_cat(), o["pattern_lh_cat"]
    denom = np.maximum(np.abs(cat_o).max(axis=1, keepdims=True), 1e-300)
    assert float(np.max(np.abs(cat_r - cat_o) / denom)) <= 1e-9
    df, ddf = ref.derv()
    assert rel(df, o["df"]) <= 1e-9 and rel(ddf, o["ddf"]) <= 1e-9
    assert rel(ref.from_buffer(), o["lnl_buf"]) <= tol
    if check_nodes:
        for slot, ls in leaf_sets(tree, o["traversal"]).items():
            part, sc = ref.fetch_branch(ls, leaf)
            assert np.array_equal(sc, o["scale"][slot]), "scale_num differs from the reference at slot %d" % slot
            got = eng.fetch_partial(slot)
            d = np.maximum(np.abs(part).reshape(part.shape[0], -1).max(axis=1), 1e-300)[:, None, None]
            assert float(np.max(np.abs(got - part) / d)) <= 1e-9
    # a second branch length on the same theta: what Newton-Raphson steps do
    L = tree.get_length(leaf, tree.adj[leaf][0])
    for t in (0.5 * L, 3.0 * L):
        ref.set_current_length(t)
        rdf = ref.derv(again=True)
        odf = eng.derv(t)
        assert rel(rdf, odf) <= 1e-9
        assert rel(ref.from_buffer(), eng.lnl_from_buffer(t)) <= tol
    ref.close()
    return lnl


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("safe", [False, True])
def test_dna_gtr_g4(oracle_lib, kernel, safe):
    tree, model, rates, props, aln = make_problem(14, 600, seed=3, gaps=0.05)
    compare(tree, aln, model, rates, props, kernel, safe)


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("safe", [False, True])
def test_protein_20_states(oracle_lib, kernel, safe):
    tree, model, rates, props, aln = make_problem(9, 150, nstates=20, seed=4, gaps=0.03)
    compare(tree, aln, model, rates, props, kernel, safe)


@pytest.mark.parametrize("kernel", [oracle.LK_AVX_FMA, oracle.LK_SSE2])
def test_mixture_and_generic_states(oracle_lib, kernel):
    # 20-state mixture (C-series shape: nmix classes x ncat rates)
    tree, model, rates, props, aln = make_problem(7, 120, nstates=20, nmix=3, seed=5)
    compare(tree, aln, model, rates, props, kernel, False)
    # generic nstates (SAFE forced by the reference): even state count
    tree, model, rates, props, aln = make_problem(8, 200, nstates=6, ncat=3, nmix=2, seed=6)
    compare(tree, aln, model, rates, props, kernel, False)


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("nstates,k", [(4, 4), (20, 3), (6, 2)])
def test_fused_mixture_rate_categories(oracle_lib, nstates, k, safe):
    """model_factory->fused_mix_rate: category c pairs mixture class c with rate category c (ncat_mix = ncat,
    tree/phylokernelnew.h:1322-1331)."""
    tree, model, rates, props, aln = make_problem(9, 180, nstates=nstates, ncat=k, nmix=k, seed=12, gaps=0.04)
    compare(tree, aln, model, rates, props, oracle.LK_AVX_FMA, safe, fused=True)


@pytest.mark.parametrize("safe", [False, True])
def test_deep_tree_scale_counts_match_reference(oracle_lib, safe):
    tree, model, rates, props, aln = make_problem(300, 120, seed=5, mean_len=0.4)
    ref = oracle.RefRun(tree, aln, model, rates, props, kernel=oracle.LK_AVX_FMA, safe=safe)
    eng = oracle.oracle_engine(aln, model, rates, props, nslots=300, safe_numeric=safe)
    lnl = ref.lnl_at_leaf(0)
    o = full_eval(eng, tree, 0)
    assert rel(lnl, o["lnl"]) <= 1e-11
    nz = 0
    for slot, ls in leaf_sets(tree, o["traversal"]).items():
        _, sc = ref.fetch_branch(ls, 0)
        assert np.array_equal(sc, o["scale"][slot])
        nz += int(sc.max() > 0)
    assert nz > 0, "the test tree must be deep enough for rescaling to fire"
    ref.close()


def test_invariant_sites(oracle_lib):
    tree, model, rates, props, aln = make_problem(10, 400, seed=8)
    compare(tree, aln, model, rates, props, oracle.LK_AVX_FMA, False, p_invar=0.25)


def test_example_phy(oracle_lib):
    import os
    here = os.path.dirname(os.path.abspath(__file__))
    aln = W.alignment_from_phylip(os.path.join(here, "golden", "example.phy"))
    tree = W.Tree.from_newick(open(os.path.join(here, "golden", "example17.nwk")).read(), aln.names)
    model = W.gtr_model()
    rates, props = W.discrete_gamma(0.5)
    for kernel in KERNELS:
        compare(tree, aln, model, rates, props, kernel, False)


def test_odd_state_count_quirk_is_the_reference_s(oracle_lib):
    """tree/phylokernelnew.h:2491 + :700-724: with odd nstates the reference's dotProductTriple counts the
    last block entry twice.  The oracle reproduces it when ref quirks are on (and the CUDA path does not,
    DESIGN.md 'deviations')."""
    tree, model, rates, props, aln = make_problem(7, 150, nstates=5, ncat=2, nmix=1, seed=4)
    ref = oracle.RefRun(tree, aln, model, rates, props, kernel=oracle.LK_AVX_FMA)
    ref.lnl_at_leaf(0)
    rdf = ref.derv()
    out = {}
    for q in (True, False):
        oracle.set_ref_quirks(q)
        eng = oracle.oracle_engine(aln, model, rates, props, nslots=8, safe_numeric=True)
        out[q] = full_eval(eng, tree, 0)
    oracle.set_ref_quirks(True)
    assert rel(rdf[0], out[True]["df"]) <= 1e-9
    ref.close()


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("nstates", [4, 20])
def test_ascertainment_bias_correction_lewis(oracle_lib, safe, nstates):
    """+ASC (Lewis): the nstates constant patterns are appended as unobserved patterns; lnL, per-pattern lnL,
    df/ddf and lnL-from-buffer carry the log(1 - prob_const) correction (tree/phylokernelnew.h:2506-2534,
    2620-2628, 2912-2930, 3160-3182)."""
    tree, model, rates, props, aln = make_problem(9, 150 if nstates == 4 else 40, nstates=nstates, seed=23)
    unobs = np.repeat(np.arange(nstates, dtype=np.uint8)[None, :], tree.ntaxa, axis=0)
    ref = oracle.RefRun(tree, aln, model, rates, props, kernel=oracle.LK_AVX_FMA, safe=safe, unobserved=unobs)
    eng = oracle.oracle_engine(aln, model, rates, props, nslots=tree.ntaxa, safe_numeric=safe, unobserved=unobs)
    leaf = 0
    c = tree.adj[leaf][0]
    tr = W.traversal(tree, c, leaf)
    L = tree.get_length(leaf, c)
    eng.update_partials(tr)
    dad, node = (LEAF, leaf), (INTERNAL, tr[-1][2])
    want = ref.lnl_at_leaf(leaf)
    got = eng.branch_lnl(dad, node, L)
    plain = oracle.oracle_engine(aln, model, rates, props, nslots=tree.ntaxa, safe_numeric=safe)
    plain.update_partials(tr)
    assert abs(plain.branch_lnl(dad, node, L) - want) > 1e-3  # the correction is not a no-op
    assert abs(got - want) <= 1e-11 * abs(want)
    assert np.allclose(eng.fetch_pattern_lh(), ref.pattern_lh(), rtol=1e-11, atol=1e-11)
    eng.build_theta(dad, node)
    df, ddf = eng.derv(L)
    rdf, rddf = ref.derv()
    assert abs(df - rdf) <= 1e-9 * max(1.0, abs(rdf)) and abs(ddf - rddf) <= 1e-9 * max(1.0, abs(rddf))
    for t in (0.5 * L, 2.0 * L):
        ref.set_current_length(t)
        assert abs(eng.lnl_from_buffer(t) - ref.from_buffer()) <= 1e-11 * abs(want)
