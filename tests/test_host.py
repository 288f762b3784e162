"""The C++ host layer (iqtree2_b200/host/): b200::PhyloTree's kernel-dispatch API, traversal builder,
Newton-Raphson glue and the PhyloSuperTree partition dispatch.

CPU tests bind the host layer to the CPU oracle through the host's test hook (same C ABI, prefix orc_) so that the
host LOGIC is covered where no GPU exists; the `gpu` tests run the same scenarios through the product library
(libb200lk.so) and compare with the oracle-backed run and with the reference-generated golden vectors."""
import os

import numpy as np
import pytest

from iqtree2_b200 import build_host, host as H
from iqtree2_b200 import workload as W
from tests.common import make_problem

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def hostlib():
    build_host.build()
    return H.lib()


@pytest.fixture()
def oracle_backend(hostlib, oracle_lib):
    H.set_kernel_library(oracle_lib, "orc_")
    yield
    H.set_kernel_library(None, None)


@pytest.fixture()
def cuda_backend(hostlib):
    H.set_kernel_library(None, None)
    yield


def host_problem(tree, model, rates, props, aln, p_invar=0.0, safe=False, device_newton=True, mem_save=None):
    # the host compresses columns itself: expand the generator's weighted patterns back into columns (first
    # appearances keep their order, so pattern p of the host alignment is pattern p of `aln`)
    cols = np.repeat(aln.states, aln.ptn_freq.astype(int), axis=1)
    ha = H.HostAlignment.from_states(cols, aln.nstates, compress=True)
    assert ha.nptn == aln.nptn
    hm = H.HostModel.from_eigen(model)
    hr = H.HostRate.categories(rates, np.asarray(props) * (1.0 - p_invar), p_invar)
    return H.HostTree(relabel(tree), ha, hm, hr, safe=safe, device_newton=device_newton, mem_save=mem_save), ha


def relabel(tree):
    """Newick with leaves named T<row> (the names HostAlignment.from_states gives its rows)."""
    saved = list(tree.names)
    tree.names = ["T%d" % i for i in range(tree.ntaxa)]
    try:
        return tree.to_newick(0, 17)
    finally:
        tree.names = saved


def oracle_eval(tree, model, rates, props, aln, leaf=0, p_invar=0.0, safe=False):
    import oracle
    from iqtree2_b200.capi import INTERNAL, LEAF
    eng = oracle.oracle_engine(aln, model, rates, props, nslots=tree.ntaxa, p_invar=p_invar, safe_numeric=safe)
    c = tree.adj[leaf][0]
    tr = W.traversal(tree, c, leaf)
    return eng.evaluate(tr, (LEAF, leaf), (INTERNAL, tr[-1][2]), tree.get_length(leaf, c))


# ---------------------------------------------------------------------------------------------------------
# host-side inputs: no kernel involved
# ---------------------------------------------------------------------------------------------------------
def test_discrete_gamma_matches_generator(hostlib):
    for alpha, ncat in [(0.5, 4), (0.7, 4), (1.0, 4), (2.5, 8), (0.05, 4), (30.0, 6)]:
        want, _ = W.discrete_gamma(alpha, ncat)
        got, props = H.HostRate.gamma(alpha, ncat).get()
        assert np.allclose(got, want, rtol=1e-9, atol=1e-13), (alpha, ncat, got, want)
        assert np.allclose(props, 1.0 / ncat)
        assert abs(got.mean() - 1.0) < 1e-12


def test_gtr_eigensystem_reconstructs_rate_matrix(hostlib):
    m = H.HostModel.gtr((1.0, 2.0, 1.5, 1.2, 2.5, 1.0), (0.3, 0.2, 0.2, 0.3))
    ev, U, Ui, fr, w = m.eigen()
    ref = W.gtr_model()
    Q = U[0] @ np.diag(ev[0]) @ Ui[0]
    assert np.allclose(Q, ref.Q[0], atol=1e-12)
    assert np.allclose(U[0] @ Ui[0], np.eye(4), atol=1e-12)
    assert np.allclose(np.sort(ev[0]), np.sort(ref.eval[0]), atol=1e-12)
    assert np.allclose(fr[0], ref.freqs[0])


def test_protein_sized_eigensystem(hostlib):
    ref = W.random_reversible_model(20, 7)
    n = 20
    pi = ref.freqs[0]
    R = ref.Q[0] / pi[None, :]
    np.fill_diagonal(R, 0.0)
    m = H.HostModel.reversible(0.5 * (R + R.T), pi)
    ev, U, Ui, _, _ = m.eigen()
    assert np.allclose(U[0] @ np.diag(ev[0]) @ Ui[0], ref.Q[0], atol=1e-11)
    assert np.allclose(U[0] @ Ui[0], np.eye(n), atol=1e-11)


def test_phylip_reader_and_pattern_compression(hostlib):
    a = H.HostAlignment.from_phylip(os.path.join(GOLD, "example.phy"))
    ref = W.alignment_from_phylip(os.path.join(GOLD, "example.phy"))
    st, fr, cc = a.patterns()
    assert a.ntaxa == ref.ntaxa and a.names() == list(ref.names)
    assert fr.sum() == ref.ptn_freq.sum()
    # same multiset of (pattern column, frequency)
    mine = sorted((bytes(st[:, p]), fr[p]) for p in range(a.nptn))
    theirs = sorted((bytes(ref.states[:, p]), ref.ptn_freq[p]) for p in range(ref.nptn))
    assert mine == theirs
    # constant patterns: const_char is the shared state
    for p in range(a.nptn):
        col = st[:, p]
        plain = col[col < 4]
        if len(plain) == len(col) and len(set(plain)) == 1:
            assert cc[p] == plain[0]
        if len(set(plain)) > 1:
            assert cc[p] > 18


def test_unbuilt_kernel_is_refused(hostlib, oracle_backend):
    tree, model, rates, props, aln = make_problem(6, 50, seed=2)
    t, _ = host_problem(tree, model, rates, props, aln)
    with pytest.raises(H.HostError, match="LK_CUDA"):
        t.set_kernel(H.LK_AVX_FMA)


def test_missing_kernel_library_fails_loudly(hostlib):
    H.set_kernel_library("/nonexistent/libb200lk.so", "b200lk_")
    try:
        tree, model, rates, props, aln = make_problem(6, 50, seed=2)
        with pytest.raises(H.HostError, match="no CPU fallback"):
            host_problem(tree, model, rates, props, aln)
    finally:
        H.set_kernel_library(None, None)


# ---------------------------------------------------------------------------------------------------------
# scenarios run against either backend
# ---------------------------------------------------------------------------------------------------------
def scenario_likelihood_every_branch(safe, nstates, nmix):
    tree, model, rates, props, aln = make_problem(9, 120 if nstates == 4 else 40, nstates=nstates, nmix=nmix, seed=11,
                                                  gaps=0.04)
    want = oracle_eval(tree, model, rates, props, aln, safe=safe)[0]
    t, _ = host_problem(tree, model, rates, props, aln, safe=safe)
    got = t.compute_likelihood()
    assert abs(got - want) <= 1e-9 * abs(want)
    # pulley principle through re-oriented slots: every branch gives the same lnL, in any visiting order
    order = list(range(t.nbranches))
    np.random.default_rng(1).shuffle(order)
    for b in order:
        assert abs(t.branch_lnl(b) - want) <= 1e-9 * abs(want)
    lnl, plh = t.compute_likelihood(want_pattern_lh=True)
    assert abs(float(np.dot(plh, aln.ptn_freq)) - lnl) <= 1e-9 * abs(lnl)
    return t


def scenario_newton(device_newton):
    """optimizeOneBranch on every branch from a perturbed start == scalar maximiser of the oracle's lnL(t)."""
    tree, model, rates, props, aln = make_problem(8, 300, seed=21)
    t, _ = host_problem(tree, model, rates, props, aln, device_newton=device_newton)
    lnl0 = t.compute_likelihood()
    lens0 = t.lengths()
    for b in range(t.nbranches):
        # derivative at the optimum vanishes and the likelihood does not decrease
        t.set_lengths(lens0)
        before = t.branch_lnl(b)
        opt = t.optimize_one_branch(b)
        after = t.from_buffer()
        assert after >= before - 1e-7
        df, ddf = t.derv(b, opt)
        if 1e-6 * 1.01 < opt < 9.5:
            assert abs(df) < 1e-3 * max(1.0, abs(ddf)) * 1e-2 + 1e-4 * abs(ddf) * 1e-6 + 1e-2
            assert ddf < 0
    t.set_lengths(lens0)
    assert abs(t.compute_likelihood() - lnl0) <= 1e-9 * abs(lnl0)
    return t


def scenario_optimize_all_branches(device_newton=True):
    tree, model, rates, props, aln = make_problem(10, 400, seed=31)
    t, _ = host_problem(tree, model, rates, props, aln, device_newton=device_newton)
    start = t.compute_likelihood()
    final = t.optimize_all_branches()
    assert final > start
    # converged: a second pass moves the likelihood by less than the tolerance, fresh evaluation agrees
    t.clear_partials()
    assert abs(t.compute_likelihood() - final) <= 1e-8 * abs(final)
    again = t.optimize_all_branches()
    assert abs(again - final) < 0.01
    return final, t.lengths()


def scenario_mem_save(nstates):
    """-mem (LM_MEM_SAVE, tree/memslot.cpp): log2(T)+1 recycled slots give the same likelihood at every branch, in
    any visiting order, and the same branch-length optimum as one slot per internal node."""
    ntaxa = 40
    tree, model, rates, props, aln = make_problem(ntaxa, 150 if nstates == 4 else 50, nstates=nstates, seed=17,
                                                  mean_len=0.15)
    want = oracle_eval(tree, model, rates, props, aln)[0]
    full, _ = host_problem(tree, model, rates, props, aln)
    assert full.max_lh_slots == ntaxa - 2
    final_full = full.optimize_all_branches()
    for frac, slots in ((0, 6), (0.25, 9), (0.5, 19)):
        t, _ = host_problem(tree, model, rates, props, aln, mem_save=frac)
        assert t.max_lh_slots == slots
        assert abs(t.compute_likelihood() - want) <= 1e-9 * abs(want)
        order = list(range(t.nbranches))
        np.random.default_rng(2).shuffle(order)
        for b in order:
            assert abs(t.branch_lnl(b) - want) <= 1e-9 * abs(want)
        final = t.optimize_all_branches()
        assert abs(final - final_full) <= 1e-7 * abs(final_full)
        assert np.allclose(t.lengths(), full.lengths(), rtol=1e-5, atol=1e-7)
    return final_full


def scenario_supertree(edge_linked, batching=True):
    tree = W.random_tree(7, 5)
    newick = relabel(tree)
    rng = np.random.default_rng(3)
    st = H.HostSuperTree(newick, edge_linked=edge_linked)
    want = []
    parts = []
    for g in range(5):
        nst = 4 if g % 2 == 0 else 20
        model = W.gtr_model() if nst == 4 else W.random_reversible_model(20, 50 + g)
        rates, props = W.discrete_gamma(0.4 + 0.2 * g)
        rate = float(rng.gamma(4.0, 0.25)) if edge_linked else 1.0
        scaled = W.random_tree(7, 5)
        for (u, v) in scaled.branches():
            scaled.set_length(u, v, tree.get_length(u, v) * rate)
        aln = W.simulate_alignment(scaled, model, rates, props, 60 + 40 * g, 100 + g, compress=True)
        want.append(oracle_eval(scaled, model, rates, props, aln)[0])
        ha = H.HostAlignment.from_states(aln.states, nst, compress=False)
        # weights of the compressed patterns are carried by repeating columns on the host side instead:
        cols = np.repeat(aln.states, aln.ptn_freq.astype(int), axis=1)
        ha = H.HostAlignment.from_states(cols, nst, compress=True)
        st.add_partition("gene%d" % g, ha, H.HostModel.from_eigen(model), H.HostRate.categories(rates, props), rate)
        parts.append((nst, aln.nptn))
    st.set_batching(batching)
    st.initialize()
    total = st.compute_likelihood()
    scores = st.part_scores()
    assert np.allclose(scores, want, rtol=1e-9)
    assert abs(total - sum(want)) <= 1e-9 * abs(sum(want))
    # cost order of computePartitionOrder: decreasing T*N*n
    order = st.part_order()
    cost = [7 * n * s for (s, n) in parts]
    assert [cost[i] for i in order] == sorted(cost, reverse=True)
    better = st.optimize_all_branches()
    assert better >= total - 1e-6
    scenario_supertree.batch_info = st.batch_info()
    return total, better


@pytest.mark.parametrize("safe,nstates,nmix", [(False, 4, 1), (True, 4, 1), (False, 20, 1), (False, 20, 2), (True, 7, 1)])
def test_host_likelihood_oracle_backend(oracle_backend, safe, nstates, nmix):
    t = scenario_likelihood_every_branch(safe, nstates, nmix)
    assert t.safe_numeric() == (safe or nstates not in (4, 20))


def test_host_newton_oracle_backend(oracle_backend):
    scenario_newton(False)


def test_host_optimize_all_branches_oracle_backend(oracle_backend):
    scenario_optimize_all_branches()


@pytest.mark.parametrize("nstates", [4, 20])
def test_host_mem_save_oracle_backend(oracle_backend, nstates):
    scenario_mem_save(nstates)


@pytest.mark.parametrize("edge_linked", [True, False])
def test_host_supertree_oracle_backend(oracle_backend, edge_linked):
    scenario_supertree(edge_linked)


def test_host_example_phy_matches_reference_golden(oracle_backend):
    """cfg1: example.phy, GTR+G4 on the committed 17-taxon tree: lnL against the value the REFERENCE's own
    FMA kernel produced (tests/golden/ref_example_phy.npz, written by make_golden.py)."""
    g = np.load(os.path.join(GOLD, "ref_example_phy.npz"), allow_pickle=True)
    ha = H.HostAlignment.from_phylip(os.path.join(GOLD, "example.phy"))
    hm = H.HostModel.gtr((1.0, 2.0, 1.5, 1.2, 2.5, 1.0), (0.3, 0.2, 0.2, 0.3))
    hr = H.HostRate.gamma(0.5, 4)
    t = H.HostTree(open(os.path.join(GOLD, "example17.nwk")).read(), ha, hm, hr)
    lnl = t.compute_likelihood()
    assert abs(lnl - float(g["lnl"])) <= 1e-9 * abs(float(g["lnl"]))
    opt = t.optimize_all_branches()
    assert opt > lnl
    check_against_reference_optimisation(t, ha, g, opt)


def check_against_reference_optimisation(t, ha, g, opt):
    """cfg1's second half: branch-length optimisation.  tests/golden/ref_example_phy.npz holds what the REFERENCE's
    own PhyloTree::optimizeAllBranches (AVX+FMA kernel, Newton-Raphson) reached from the same start: the final lnL
    and every optimised branch length, keyed by the leaf set behind the branch."""
    assert abs(opt - float(g["opt_lnl"])) <= 1e-7 * abs(float(g["opt_lnl"]))
    names = [str(x) for x in g["names"]]
    row_of = {nm: i for i, nm in enumerate(ha.names())}
    taxon_row = [row_of[nm] for nm in names]                 # reference taxon id -> host alignment row
    want = {}
    for key, ln in zip(g["opt_splits"], g["opt_lengths"]):
        want[frozenset(taxon_row[int(x)] for x in str(key).split(","))] = float(ln)
    everything = frozenset(range(ha.ntaxa))
    lens = t.lengths()
    matched = 0
    for b in range(t.nbranches):
        side = t.branch_split(b)
        ref_len = want.get(side, want.get(everything - side))
        assert ref_len is not None
        assert abs(lens[b] - ref_len) <= 1e-5 * max(ref_len, 1e-3), (b, lens[b], ref_len)
        matched += 1
    assert matched == 2 * ha.ntaxa - 3


# ---------------------------------------------------------------------------------------------------------
# the same scenarios through the product library on the GPU
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("safe,nstates,nmix", [(False, 4, 1), (True, 4, 1), (False, 20, 1), (False, 20, 2), (True, 7, 1)])
def test_host_likelihood_cuda(cuda_backend, safe, nstates, nmix):
    scenario_likelihood_every_branch(safe, nstates, nmix)


@pytest.mark.gpu
@pytest.mark.parametrize("device_newton", [False, True])
def test_host_newton_cuda(cuda_backend, device_newton):
    scenario_newton(device_newton)


@pytest.mark.gpu
def test_host_device_newton_equals_host_newton(cuda_backend, oracle_lib):
    a, la = scenario_optimize_all_branches(device_newton=True)
    b, lb = scenario_optimize_all_branches(device_newton=False)
    assert abs(a - b) <= 1e-9 * abs(b)
    assert np.allclose(la, lb, rtol=1e-7, atol=1e-9)
    H.set_kernel_library(oracle_lib, "orc_")
    try:
        c, lc = scenario_optimize_all_branches()
    finally:
        H.set_kernel_library(None, None)
    assert abs(a - c) <= 1e-9 * abs(c)
    assert np.allclose(la, lc, rtol=1e-6, atol=1e-8)


@pytest.mark.gpu
@pytest.mark.parametrize("nstates", [4, 20, 7])
def test_host_mem_save_cuda(cuda_backend, nstates):
    scenario_mem_save(nstates)


@pytest.mark.gpu
@pytest.mark.parametrize("edge_linked", [True, False])
def test_host_supertree_cuda(cuda_backend, oracle_lib, edge_linked):
    total, better = scenario_supertree(edge_linked)
    # the three DNA genes and the two protein genes each ran as one batch of launches
    nbatches, nbatched, calls = scenario_supertree.batch_info
    assert (nbatches, nbatched) == (2, 5) and calls > 0
    t1, b1 = scenario_supertree(edge_linked, batching=False)
    assert scenario_supertree.batch_info[0] == 0
    assert abs(total - t1) <= 1e-10 * abs(t1) and abs(better - b1) <= 1e-8 * abs(b1)
    H.set_kernel_library(oracle_lib, "orc_")
    try:
        t2, b2 = scenario_supertree(edge_linked)
    finally:
        H.set_kernel_library(None, None)
    assert abs(total - t2) <= 1e-9 * abs(t2)
    assert abs(better - b2) <= 1e-7 * abs(b2)


@pytest.mark.gpu
def test_host_example_phy_cuda(cuda_backend):
    g = np.load(os.path.join(GOLD, "ref_example_phy.npz"), allow_pickle=True)
    ha = H.HostAlignment.from_phylip(os.path.join(GOLD, "example.phy"))
    hm = H.HostModel.gtr((1.0, 2.0, 1.5, 1.2, 2.5, 1.0), (0.3, 0.2, 0.2, 0.3))
    t = H.HostTree(open(os.path.join(GOLD, "example17.nwk")).read(), ha, hm, H.HostRate.gamma(0.5, 4))
    lnl = t.compute_likelihood()
    assert abs(lnl - float(g["lnl"])) <= 1e-9 * abs(float(g["lnl"]))
    opt = t.optimize_all_branches()
    check_against_reference_optimisation(t, ha, g, opt)


def scenario_asc():
    """+ASC through the host layer: Params::asc_lewis appends the constant patterns; Newton-Raphson falls back to
    the host-driven iteration where the device one does not offer the correction."""
    import oracle
    from iqtree2_b200.capi import INTERNAL, LEAF
    tree, model, rates, props, aln = make_problem(8, 200, seed=41)
    unobs = np.repeat(np.arange(4, dtype=np.uint8)[None, :], tree.ntaxa, axis=0)
    eng = oracle.oracle_engine(aln, model, rates, props, nslots=tree.ntaxa, unobserved=unobs)
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    want = eng.evaluate(tr, (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c))[0]
    cols = np.repeat(aln.states, aln.ptn_freq.astype(int), axis=1)
    ha = H.HostAlignment.from_states(cols, 4, compress=True)
    t = H.HostTree(relabel(tree), ha, H.HostModel.from_eigen(model), H.HostRate.categories(rates, props), asc=True)
    got = t.compute_likelihood()
    assert abs(got - want) <= 1e-9 * abs(want)
    better = t.optimize_all_branches()
    assert better >= got
    return got, better


def test_host_asc_oracle_backend(oracle_backend):
    scenario_asc()


@pytest.mark.gpu
def test_host_asc_cuda(cuda_backend, oracle_lib):
    a = scenario_asc()
    H.set_kernel_library(oracle_lib, "orc_")
    try:
        b = scenario_asc()
    finally:
        H.set_kernel_library(None, None)
    assert abs(a[0] - b[0]) <= 1e-9 * abs(b[0]) and abs(a[1] - b[1]) <= 1e-7 * abs(b[1])
