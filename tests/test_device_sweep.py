"""Branch lengths resident on the device (sweep.cu; SURVEY.md 8f rank 2, second half): b200lk_set_lengths /
b200lk_update_partials_slots / b200lk_optimize_branch_slot / b200lk_get_lengths through the C ABI, and
b200::PhyloTree::optimizeAllBranches enqueuing a whole sweep with ONE read-back of the length vector -- against the
branch-by-branch sweep (tree/phylotree.cpp:2689-2765 -> :2585-2635) and the reference's own optimum."""
import numpy as np
import pytest

import oracle
from iqtree2_b200 import build_host, host as H, workload as W
from iqtree2_b200.capi import INTERNAL, LEAF, engine_for
from tests.common import make_problem, rel

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nstates,nsite,safe", [(4, 2500, False), (4, 600, True), (20, 300, False), (7, 200, True)])
def test_slot_entry_points_equal_the_ordinary_ones(nstates, nsite, safe):
    tree, model, rates, props, aln = make_problem(14, nsite, nstates=nstates, seed=71, gaps=0.04)
    a = engine_for(aln, model, rates, props, nslots=14, safe_numeric=safe)
    b = engine_for(aln, model, rates, props, nslots=14, safe_numeric=safe)
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    dad, node, L = (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c)
    # a length table indexed by the position of the child in the packed list, plus one slot for the root branch
    lens = [ch[2] for e in tr for ch in e[3]] + [L]
    b.set_lengths(lens)
    slots = list(range(len(lens) - 1))
    junk = [(v, u, s, [(k, i, 123.0) for k, i, _ in ch]) for v, u, s, ch in tr]  # the lengths in the entries must be ignored
    a.update_partials(tr)
    b.update_partials_slots(junk, slots)
    for (_, _, s, _) in tr:
        assert np.array_equal(a.fetch_scale_num(s), b.fetch_scale_num(s))
        pa, pb = a.fetch_partial(s), b.fetch_partial(s)
        assert rel(pa, pb) <= 1e-12  # exp() of the host against exp() of the device in the 4-state factors
    want = a.optimize_branch(dad, node, L)
    b.optimize_branch_slot(dad, node, len(lens) - 1)
    got, aux = b.get_lengths()
    assert abs(got[-1] - want[0]) <= 1e-9 and int(aux[-1][0]) == want[2] and aux[-1][1] == 0.0
    assert np.array_equal(got[:-1], np.asarray(lens[:-1]))  # only the optimised slot changed
    assert rel(b.lnl_from_buffer(got[-1]), a.lnl_from_buffer(want[0])) <= 1e-9  # theta valid as after optimize_branch
    # mixed: some children by slot, some by value
    half = [s if j % 2 == 0 else -1 for j, s in enumerate(slots)]
    b.update_partials_slots(tr, half)
    assert rel(b.fetch_partial(tr[-1][2]), a.fetch_partial(tr[-1][2])) <= 1e-12


def _host_tree(tree, model, rates, props, aln, **kw):
    from tests.test_host import relabel
    cols = np.repeat(aln.states, aln.ptn_freq.astype(int), axis=1)
    ha = H.HostAlignment.from_states(cols, aln.nstates, compress=True)
    return H.HostTree(relabel(tree), ha, H.HostModel.from_eigen(model), H.HostRate.categories(rates, props), **kw)


@pytest.mark.parametrize("nstates,ntaxa,nsite,safe", [(4, 30, 2000, False), (4, 12, 300, True), (20, 10, 200, False), (5, 9, 250, False)])
def test_device_sweep_equals_branch_by_branch_sweep(nstates, ntaxa, nsite, safe):
    build_host.build()
    H.set_kernel_library(None, None)
    tree, model, rates, props, aln = make_problem(ntaxa, nsite, nstates=nstates, seed=73)
    one = _host_tree(tree, model, rates, props, aln, safe=safe)
    one.set_device_sweep(False)
    dev = _host_tree(tree, model, rates, props, aln, safe=safe)
    l1, l2 = one.optimize_all_branches(iterations=1), dev.optimize_all_branches(iterations=1)
    assert one.device_sweeps() == 0 and dev.device_sweeps() == 1
    assert rel(l1, l2) <= 1e-9 and np.max(np.abs(one.lengths() - dev.lengths())) <= 1e-8
    # far fewer blocking calls: the sweep reads back once, not once per branch
    f1, f2 = one.optimize_all_branches(), dev.optimize_all_branches()
    assert rel(f1, f2) <= 1e-9 and np.max(np.abs(one.lengths() - dev.lengths())) <= 1e-6
    assert dev.device_sweeps() >= 2


def test_device_sweep_falls_back_when_a_length_runs_to_the_upper_bound():
    """optimizeOneBranch compares an optimum above 0.95 max_branch_length with the old length (tree/phylotree.cpp:2607-2616);
    a device sweep cannot, so it gives up and the sweep is repeated branch by branch -- same result as without it."""
    build_host.build()
    H.set_kernel_library(None, None)
    tree, model, rates, props, aln = make_problem(8, 400, seed=75)
    res = []
    for sweep in (False, True):
        t = _host_tree(tree, model, rates, props, aln)
        t.set_device_sweep(sweep)
        t.set_branch_bounds(1e-6, 0.06)  # most optima lie above 0.95 * 0.06
        res.append((t.optimize_all_branches(iterations=2), t.lengths(), t.device_sweeps()))
    assert res[1][2] == 0  # every device sweep was given up
    assert rel(res[0][0], res[1][0]) <= 1e-9 and np.max(np.abs(res[0][1] - res[1][1])) <= 1e-9
