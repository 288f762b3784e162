"""Host mirror (iqtree2_b200/host/) of the callers of the a15 / +ASC / robust variants: b200::PhyloTreeMixlen
(tree/phylotreemixlen.cpp: per-class lengths through the dispatch members, computeFuncDervMulti, the EM branch
update), Params::asc_holder (Alignment::getUnobservedConstPatterns for ASC_VARIANT_MISSING), Params::robust_phy_keep /
robust_median, and PhyloTree::computePatternLhCat (tree/phylotree.cpp:1330-1388).

CPU tests bind the host to the CPU oracle (same C ABI under prefix orc_) to cover the host LOGIC; `gpu` tests run the
same scenarios through libb200lk.so and compare with the oracle-backed run."""
import numpy as np
import pytest

import oracle
from iqtree2_b200 import host as H
from iqtree2_b200 import workload as W
from iqtree2_b200.capi import INTERNAL, LEAF
from tests.common import make_problem, rel
from tests.test_host import cuda_backend, hostlib, oracle_backend, relabel  # noqa: F401  (fixtures)
from tests.test_mixlen_holder import holder_problem, holder_unobserved, mix_child_len, mixlen_problem


def host_tree(tree, model, rates, props, aln, **kw):
    cols = np.repeat(aln.states, aln.ptn_freq.astype(int), axis=1)
    ha = H.HostAlignment.from_states(cols, aln.nstates, compress=True)
    assert ha.nptn == aln.nptn
    return H.HostTree(relabel(tree), ha, H.HostModel.from_eigen(model), H.HostRate.categories(rates, props), **kw)


def engine_lnl(eng, tree, leaf=0):
    c = tree.adj[leaf][0]
    tr = W.traversal(tree, c, leaf)
    eng.update_partials(tr)
    return eng.branch_lnl((LEAF, leaf), (INTERNAL, tr[-1][2]), tree.get_length(leaf, c))


# ------------------------------------------------------------------------------------------------ scenarios
def scenario_holder(backend_is_cuda):
    tree, model, rates, props, aln = holder_problem(4, seed=205)
    t = host_tree(tree, model, rates, props, aln, asc_holder=True)
    lnl = t.compute_likelihood()
    # the same through the bare C ABI with the patterns built by the test
    eng = oracle.oracle_engine(aln, model, rates, props, nslots=tree.ntaxa, unobserved=holder_unobserved(aln),
                               asc_holder=True)
    want = engine_lnl(eng, tree)
    assert rel(lnl, want) <= 1e-9
    opt = t.optimize_all_branches(3, 0.01)  # host-driven Newton-Raphson (the device iteration refuses +ASC)
    assert opt > lnl
    return lnl, opt, t.lengths()


def scenario_robust():
    tree, model, rates, props, aln = make_problem(10, 400, nstates=4, seed=207)
    out = []
    for rb in ((0.8, False), (1.0, True)):
        t = host_tree(tree, model, rates, props, aln, robust=rb)
        eng = oracle.oracle_engine(aln, model, rates, props, nslots=tree.ntaxa)
        eng.set_robust(*rb)
        got, want = t.compute_likelihood(), engine_lnl(eng, tree)
        assert rel(got, want) <= 1e-9
        out.append(got)
    return out


def scenario_pattern_lh_cat():
    tree, model, rates, props, aln = make_problem(8, 200, nstates=4, ncat=4, nmix=3, seed=209)
    t = host_tree(tree, model, rates, props, aln)
    lnl, full = t.pattern_lh_cat(4)          # WSL_MIXTURE_RATECAT: untouched [nptn][nmix * ncat], mixture-major
    assert full.shape == (aln.nptn, 12)
    l2, per_mix = t.pattern_lh_cat(3)        # WSL_MIXTURE
    l3, per_rate = t.pattern_lh_cat(2)       # WSL_RATECAT
    assert per_mix.shape == (aln.nptn, 3) and per_rate.shape == (aln.nptn, 4)
    cube = full.reshape(aln.nptn, 3, 4)
    assert np.allclose(per_mix, cube.sum(axis=2), rtol=1e-13) and np.allclose(per_rate, cube.sum(axis=1), rtol=1e-13)
    assert lnl == l2 == l3
    # the pattern log-likelihoods are the log of the row sums (no +I, no rescaling at this size)
    _, plh = t.compute_likelihood(want_pattern_lh=True)
    assert np.allclose(np.log(full.sum(axis=1)), plh, rtol=1e-10)
    return lnl, full


def scenario_mixlen():
    tree, model, rates, props, aln, mult = mixlen_problem(4, 3, seed=211, ntaxa=10, nsite=500)
    t = host_tree(tree, model, rates, props, aln, mixlen=mult)
    lnl = t.compute_likelihood()
    eng = oracle.oracle_engine(aln, model, rates, props, nslots=tree.ntaxa)
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    eng.update_partials_mixlen(tr, mix_child_len(tr, mult))
    lens0 = tree.get_length(0, c) * mult
    want = eng.branch_lnl_mixlen((LEAF, 0), (INTERNAL, tr[-1][2]), lens0)
    assert rel(lnl, want) <= 1e-9
    # computeFuncDervMulti at some branch: gradient by central differences of the lnL
    bid = 3
    lens = t.mix_lengths()[bid].copy()
    df, ddf, l0 = t.derv_multi(bid, lens)
    assert rel(l0, lnl) <= 1e-9 and np.allclose(ddf, ddf.T, rtol=1e-9, atol=1e-9)
    for k in range(3):
        h = 1e-6
        up, dn = lens.copy(), lens.copy()
        up[k] += h
        dn[k] -= h
        num = (t.derv_multi(bid, up)[2] - t.derv_multi(bid, dn)[2]) / (2 * h)
        assert abs(num - df[k]) <= 1e-4 * max(1.0, abs(df[k])), (k, num, df[k])
    t.derv_multi(bid, lens)
    # EM sweeps over all branches: the likelihood must not decrease and the class lengths must move apart
    opt = t.optimize_all_branches(2, 0.01)
    assert opt >= lnl - 1e-6
    return lnl, df, ddf, opt, t.mix_lengths()


# ------------------------------------------------------------------------------------------------ CPU (oracle-backed)
def test_host_holder_asc(oracle_backend):
    scenario_holder(False)


def test_host_robust(oracle_backend):
    scenario_robust()


def test_host_pattern_lh_cat(oracle_backend):
    scenario_pattern_lh_cat()


def test_host_mixlen(oracle_backend):
    scenario_mixlen()


# ------------------------------------------------------------------------------------------------ GPU (libb200lk.so)
def _both(fn, *a):
    H.set_kernel_library(oracle.ORACLE_LIB, "orc_")
    want = fn(*a)
    H.set_kernel_library(None, None)
    got = fn(*a)
    return got, want


def _close(g, w, tol=1e-9):
    if isinstance(g, (tuple, list)):
        assert len(g) == len(w)
        for x, y in zip(g, w):
            _close(x, y, tol)
    else:
        assert rel(g, w) <= tol, (g, w)


@pytest.mark.gpu
def test_host_holder_asc_cuda(hostlib):
    g, w = _both(scenario_holder, True)
    _close(g[:2], w[:2], 1e-8)
    assert np.allclose(g[2], w[2], rtol=1e-5, atol=1e-7)


@pytest.mark.gpu
def test_host_robust_cuda(hostlib):
    _close(*_both(scenario_robust))


@pytest.mark.gpu
def test_host_pattern_lh_cat_cuda(hostlib):
    g, w = _both(scenario_pattern_lh_cat)
    _close(g[0], w[0])
    assert np.allclose(g[1], w[1], rtol=1e-9, atol=1e-300)


@pytest.mark.gpu
def test_host_mixlen_cuda(hostlib):
    g, w = _both(scenario_mixlen)
    _close(g[:3], w[:3])
    _close(g[3], w[3], 1e-7)
    assert np.allclose(g[4], w[4], rtol=1e-4, atol=1e-6)
