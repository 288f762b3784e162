"""Non-reversible kernel family (tree/phylokernelnonrev.h): the CPU restatement
(oracle/nonrev_oracle.inc) is pinned against the reference's own computeNonrev* kernels (oracle/_ref), and the CUDA
path is compared with the restatement through the C ABI."""
import numpy as np
import pytest

import oracle
from iqtree2_b200 import workload as W
from tests.common import _RefModelShim, nonrev_engine, nonrev_eval, rel

needs_ref = pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref not built (needs /root/reference)")


def rooted_problem(ntaxa, nsite, nstates=4, ncat=4, nmix=1, seed=3, mean_len=0.1, gaps=0.05):
    """A rooted tree (root = the LAST leaf of `tree`), a non-reversible model, data simulated on the same topology."""
    tree = W.random_tree(ntaxa + 1, seed, mean_len=mean_len)
    nm = W.random_nonrev_model(nstates, seed + 50, nmix)
    rates, props = W.discrete_gamma(0.7, ncat)
    sim_tree = W.random_tree(ntaxa, seed + 1, mean_len=mean_len)
    sim_model = W.gtr_model() if nstates == 4 else W.random_reversible_model(nstates, seed + 9)
    aln = W.simulate_alignment(sim_tree, sim_model, rates, props, nsite, seed + 7)
    if gaps > 0:
        rng = np.random.default_rng(seed + 13)
        mask = rng.random(aln.states.shape) < gaps
        amb = rng.integers(nstates, aln.state_unknown + 1, size=aln.states.shape).astype(np.uint8)
        aln.states[mask] = amb[mask]
    return tree, nm, rates, props, aln


def oracle_nonrev(aln, nm, rates, props, **kw):
    if not __import__("os").path.exists(oracle.ORACLE_LIB):
        oracle.build(with_ref=False)
    return nonrev_engine(aln, nm, rates, props, lib_path=oracle.ORACLE_LIB, prefix="orc_", **kw)


def leaf_set(tree, v, u):
    """leaves of the subtree at v seen from u"""
    out, stack = [], [(v, u)]
    while stack:
        x, y = stack.pop()
        if tree.is_leaf(x):
            out.append(x)
        else:
            stack += [(w, x) for w in tree.adj[x] if w != y]
    return sorted(out)


def internal_branches(tree):
    return [(a, b) for (a, b) in tree.branches() if not tree.is_leaf(a) and not tree.is_leaf(b)]


@needs_ref
@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("nstates,ncat,nmix", [(4, 4, 1), (4, 1, 1), (20, 2, 1), (6, 3, 2), (4, 2, 2)])
def test_oracle_vs_reference_rooted(nstates, ncat, nmix, safe):
    ntaxa = 9
    tree, nm, rates, props, aln = rooted_problem(ntaxa, 300, nstates, ncat, nmix)
    root = ntaxa
    safe = safe or nstates not in (4, 20)   # setLikelihoodKernel forces SAFE_LH for other state counts (tree/phylotreesse.cpp:95-96)
    ref = oracle.RefRun(tree, aln, _RefModelShim(nm), rates, props, safe=safe, nonrev=(1, nm.Q, nm.ceval, nm.cevec, nm.cinv_evec))
    eng = oracle_nonrev(aln, nm, rates, props, nslots=4 * ntaxa, rooted=True, safe_numeric=safe)
    assert ref.safe_numeric == safe
    # the root branch, two leaf branches, every internal branch: the same lnL, the reference's derivatives
    cases = [(root, tree.adj[root][0]), (0, tree.adj[0][0]), (ntaxa - 1, tree.adj[ntaxa - 1][0])] + internal_branches(tree)
    lnls = []
    for a, b in cases:
        if tree.is_leaf(a):
            want = ref.lnl_at_leaf(a)
        else:
            want = ref.lnl_at_branch(leaf_set(tree, b, a), from_taxon=leaf_set(tree, a, b)[0])
        got = nonrev_eval(eng, tree, a, b, root)
        assert rel(got["lnl"], want) <= 1e-10, (a, b, got["lnl"], want)
        assert rel(got["pattern_lh"], ref.pattern_lh()) <= 1e-10
        g, w = got["pattern_lh_cat"], ref.pattern_lh_cat()
        assert float(np.max(np.abs(g - w) / np.maximum(np.abs(w).max(axis=1, keepdims=True), 1e-300))) <= 1e-9
        if a != root:
            df, ddf = ref.derv()
            assert rel(got["df"], df) <= 1e-9 and rel(got["ddf"], ddf) <= 1e-9, (a, b, got["df"], df, got["ddf"], ddf)
            L = tree.get_length(a, b)
            ref.set_current_length(L * 1.3)
            assert rel(got["lnl_buf"], ref.from_buffer()) <= 1e-10
            ref.set_current_length(L)
        lnls.append(got["lnl"])
    assert max(lnls) - min(lnls) <= 1e-8 * abs(lnls[0])   # the likelihood does not depend on the branch it is taken at
    ref.close()
    eng.close()
