"""b200lk_create_sharded (include/b200lk.h; SURVEY.md 8b `device_mask`, 8e): one problem split by site pattern over
several devices of one process, every entry point fanning out and summing in device-index order.  Checked against the
single-device context (and the oracle) through the C ABI and through the C++ host layer (Params::devices).  On a box
with one GPU the shards share device 0 (an ordinal may repeat); with two or more, real devices are used as well."""
import numpy as np
import pytest

import oracle
from iqtree2_b200 import build_host, host as H, workload as W
from iqtree2_b200.capi import INTERNAL, LEAF, B200LKError, Engine, engine_for
from tests.common import assert_parity, full_eval, make_problem, rel

pytestmark = pytest.mark.gpu
TOL = 1e-9


def device_sets():
    import torch
    n = torch.cuda.device_count()
    sets = [[0, 0], [0, 0, 0]]
    if n >= 2:
        sets.append([0, 1])
    if n >= 4:
        sets.append([3, 1, 0, 2])
    return sets


@pytest.mark.parametrize("nstates,nmix,ntaxa,nsite,safe", [
    (4, 1, 14, 900, False), (4, 1, 14, 900, True), (20, 1, 9, 300, False), (20, 3, 6, 150, False), (5, 1, 8, 200, True),
])
def test_sharded_context_equals_single_device_and_oracle(oracle_lib, nstates, nmix, ntaxa, nsite, safe):
    tree, model, rates, props, aln = make_problem(ntaxa, nsite, nstates=nstates, nmix=nmix, seed=17, gaps=0.04)
    oracle.set_ref_quirks(False)
    o = oracle.oracle_engine(aln, model, rates, props, nslots=ntaxa, safe_numeric=safe)
    want = full_eval(o, tree, 0)
    one = full_eval(engine_for(aln, model, rates, props, nslots=ntaxa, safe_numeric=safe), tree, 0)
    for devs in device_sets():
        g = engine_for(aln, model, rates, props, nslots=ntaxa, safe_numeric=safe, devices=devs)
        info = g.shard_info()
        assert [d for d, _, _ in info] == devs
        assert info[0][1] == 0 and info[-1][2] == aln.nptn and all(a[2] == b[1] for a, b in zip(info, info[1:]))
        assert all(b % 32 == 0 or b == aln.nptn for _, b, _ in info)  # whole 32-pattern tiles
        got = full_eval(g, tree, 0)
        assert_parity(got, want, TOL)
        # per-pattern data gathered in pattern order, identical to the single-device context
        for k in ("pattern_lh", "pattern_lh_cat", "theta", "bscale", "partial_top"):
            assert np.array_equal(np.asarray(got[k]), np.asarray(one[k])), k
        # a second evaluation gives the same bits (ordered host sum)
        again = full_eval(g, tree, 0)
        assert again["lnl"] == got["lnl"] and again["df"] == got["df"] and again["ddf"] == got["ddf"]
        g.close()


def test_sharded_one_call_evaluation_newton_and_rell(oracle_lib):
    tree, model, rates, props, aln = make_problem(12, 700, seed=23)
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    dad, node, L = (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c)
    one = engine_for(aln, model, rates, props, nslots=12)
    want = one.evaluate(tr, dad, node, L)
    wopt = one.optimize_branch(dad, node, L)
    rng = np.random.default_rng(5)
    boot = rng.multinomial(int(aln.ptn_freq.sum()), aln.ptn_freq / aln.ptn_freq.sum(), size=7).astype(np.float32)
    one.upload_boot_samples(boot)
    one.lnl_from_buffer(wopt[0])
    wrell = one.rell()
    for devs in device_sets():
        g = engine_for(aln, model, rates, props, nslots=12, devices=devs)
        assert rel(g.evaluate(tr, dad, node, L), want) <= TOL
        gopt = g.optimize_branch(dad, node, L)
        assert abs(gopt[0] - wopt[0]) <= 1e-9 and gopt[2] == wopt[2]  # same optimum, same number of evaluations
        g.upload_boot_samples(boot)
        g.lnl_from_buffer(gopt[0])
        assert rel(g.rell(), wrell) <= 1e-9
        upd, launches = g.counters()
        assert upd > 0 and launches > 0 and g.min_traffic() > 0
        g.close()


def test_sharded_more_shards_than_tiles_and_refusals():
    tree, model, rates, props, aln = make_problem(6, 40, seed=3)  # < 32 patterns after compression: one tile
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    one = engine_for(aln, model, rates, props, nslots=6)
    g = engine_for(aln, model, rates, props, nslots=6, devices=[0, 0, 0])
    assert sum(1 for _, b, e in g.shard_info() if e > b) == (aln.nptn + 31) // 32
    args = (tr, (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c))
    assert rel(g.evaluate(*args), one.evaluate(*args)) <= TOL
    assert np.array_equal(g.fetch_pattern_lh(), one.fetch_pattern_lh())
    with pytest.raises(B200LKError, match="-5"):
        g.set_stream(0)
    with pytest.raises(B200LKError, match="-5"):
        g.lnl_from_buffer_noalloc(0.1)
    with pytest.raises(B200LKError, match="-5"):  # +ASC cannot be split by pattern
        Engine(ntaxa=6, nstates=4, ncat=4, nptn=40, nslots=6, n_state_codes=19, n_unobserved=4, devices=[0, 0])


def _host_tree(tree, model, rates, props, aln, devices):
    from tests.test_host import relabel
    cols = np.repeat(aln.states, aln.ptn_freq.astype(int), axis=1)
    ha = H.HostAlignment.from_states(cols, aln.nstates, compress=True)
    return H.HostTree(relabel(tree), ha, H.HostModel.from_eigen(model), H.HostRate.categories(rates, props), devices=devices)


@pytest.mark.parametrize("nstates", [4, 20])
def test_host_layer_optimises_all_branches_on_several_devices(nstates):
    """b200::PhyloTree with Params::devices runs optimizeAllBranches over pattern shards through
    libb200host.so and reaches the same optimum as on one device."""
    build_host.build()
    H.set_kernel_library(None, None)
    tree, model, rates, props, aln = make_problem(10, 500 if nstates == 4 else 160, nstates=nstates, seed=29)
    t1 = _host_tree(tree, model, rates, props, aln, None)
    l1 = t1.compute_likelihood()
    o1 = t1.optimize_all_branches()
    len1 = t1.lengths()
    for devs in device_sets():
        tn = _host_tree(tree, model, rates, props, aln, devs)
        ln, plh = tn.compute_likelihood(want_pattern_lh=True)
        assert rel(ln, l1) <= TOL and abs(float(np.dot(plh, tn.aln.patterns()[1])) - ln) <= 1e-9 * abs(ln)
        on = tn.optimize_all_branches()
        assert rel(on, o1) <= 1e-9
        assert np.max(np.abs(tn.lengths() - len1)) <= 1e-6
        tn.close()
    t1.close()
