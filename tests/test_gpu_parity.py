"""GPU parity tests: the CUDA path through the C ABI against the CPU oracle on the same
seeded inputs.  Bar (BASELINE.json north_star): scale counts bit-exact; per-pattern and
total lnL, df, ddf within 1e-9 relative in fp64."""
import numpy as np
import pytest

import oracle
from iqtree2_b200 import workload as W
from iqtree2_b200.capi import INTERNAL, LEAF, engine_for
from tests.common import assert_parity, full_eval, make_problem, rel

pytestmark = pytest.mark.gpu
TOL = 1e-9


def both(aln, model, rates, props, **kw):
    oracle.set_ref_quirks(False)  # odd nstates: the CUDA path computes the true derivative
    return engine_for(aln, model, rates, props, **kw), oracle.oracle_engine(aln, model, rates, props, **kw)


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("nstates,ncat,nmix,ntaxa,nsite", [
    (4, 4, 1, 12, 700), (4, 1, 1, 7, 301), (4, 3, 1, 9, 333), (4, 2, 2, 11, 200), (4, 2, 1, 40, 150), (20, 4, 1, 10, 300), (20, 4, 3, 6, 120), (5, 3, 2, 9, 257),
    (61, 2, 1, 5, 40),
])
def test_full_evaluation_matches_oracle(oracle_lib, safe, nstates, ncat, nmix, ntaxa, nsite):
    tree, model, rates, props, aln = make_problem(ntaxa, nsite, nstates=nstates, ncat=ncat, nmix=nmix, seed=21,
                                                  gaps=0.05)
    g, o = both(aln, model, rates, props, nslots=ntaxa, safe_numeric=safe)
    for leaf in (0, ntaxa - 1):
        assert_parity(full_eval(g, tree, leaf), full_eval(o, tree, leaf), TOL)


@pytest.mark.parametrize("safe", [False, True])
def test_scaling_counts_bit_exact_on_deep_tree(oracle_lib, safe):
    tree, model, rates, props, aln = make_problem(300, 400, seed=5, mean_len=0.4)
    g, o = both(aln, model, rates, props, nslots=300, safe_numeric=safe)
    rg, ro = full_eval(g, tree, 0), full_eval(o, tree, 0)
    assert max(int(s.max()) for s in ro["scale"].values()) >= 1
    assert_parity(rg, ro, TOL)


def test_example_phy_fixture(oracle_lib):
    import os
    here = os.path.dirname(os.path.abspath(__file__))
    aln = W.alignment_from_phylip(os.path.join(here, "golden", "example.phy"))
    tree = W.Tree.from_newick(open(os.path.join(here, "golden", "example17.nwk")).read(), aln.names)
    model = W.gtr_model()
    rates, props = W.discrete_gamma(0.5)
    g, o = both(aln, model, rates, props, nslots=20)
    assert_parity(full_eval(g, tree, 0), full_eval(o, tree, 0), TOL)


def test_internal_branch_and_newton_steps(oracle_lib):
    tree, model, rates, props, aln = make_problem(14, 500, seed=8)
    g, o = both(aln, model, rates, props, nslots=40)
    u = tree.ntaxa
    v = [w for w in tree.adj[u] if not tree.is_leaf(w)][0]
    t1 = W.traversal(tree, u, v)
    t2 = [(a, b, s + 20, [(k, (i + 20 if k == INTERNAL else i), l) for k, i, l in ch])
          for a, b, s, ch in W.traversal(tree, v, u)]
    for e in (g, o):
        e.update_partials(t1)
        e.update_partials(t2)
    dad, node = (INTERNAL, t2[-1][2]), (INTERNAL, t1[-1][2])
    L = tree.get_length(u, v)
    assert rel(g.branch_lnl(dad, node, L), o.branch_lnl(dad, node, L)) <= TOL
    g.build_theta(dad, node)
    o.build_theta(dad, node)
    for t in (L, 0.5 * L, 2 * L, 1e-6, 5.0):
        assert rel(g.derv(t), o.derv(t)) <= TOL
        assert rel(g.lnl_from_buffer(t), o.lnl_from_buffer(t)) <= TOL


def test_multifurcating_node(oracle_lib):
    names = list("ABCDEFGH")
    t = W.Tree.from_newick("(A:0.1,B:0.2,(C:0.1,D:0.3,(E:0.2,F:0.1):0.05,G:0.3,H:0.02):0.15);", names)
    model = W.gtr_model()
    rates, props = W.discrete_gamma(0.7)
    aln = W.simulate_alignment(W.random_tree(8, 4), model, rates, props, 333, 11, compress=True)
    aln.names = names
    for safe in (False, True):
        g, o = both(aln, model, rates, props, nslots=8, safe_numeric=safe)
        assert_parity(full_eval(g, t, 0), full_eval(o, t, 0), TOL)


def test_evaluate_one_call_and_edge_sizes(oracle_lib):
    for nsite in (1, 2, 31, 129):
        tree, model, rates, props, aln = make_problem(5, nsite, seed=nsite, compress=False)
        g, o = both(aln, model, rates, props, nslots=5)
        c = tree.adj[0][0]
        tr = W.traversal(tree, c, 0)
        a = g.evaluate(tr, (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c))
        b = o.evaluate(tr, (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c))
        assert rel(a, b) <= TOL


def test_large_problem_properties():
    """Full-size-ish run checked through size-independent properties: lnL is invariant under
    re-rooting, equals sum f*pattern_lh, equals lnL-from-buffer, and df matches a finite difference."""
    tree, model, rates, props, aln = make_problem(60, 200000, seed=77, compress=False)
    g = engine_for(aln, model, rates, props, nslots=60)
    vals = []
    for leaf in (0, 17):
        r = full_eval(g, tree, leaf)
        vals.append(r["lnl"])
        assert rel(r["lnl"], r["lnl_buf"]) < 1e-11
        assert rel(r["lnl"], float((r["pattern_lh"] * aln.ptn_freq).sum())) < 1e-11
        L = tree.get_length(leaf, tree.adj[leaf][0])
        h = 1e-5
        fd = (g.lnl_from_buffer(L + h) - g.lnl_from_buffer(L - h)) / (2 * h)
        assert abs(fd - r["df"]) <= 1e-5 * max(1.0, abs(r["df"]))
    assert abs(vals[0] - vals[1]) <= 1e-10 * abs(vals[0])


@pytest.mark.parametrize("env", [{"B200LK_LEVEL_BATCHED": "1"}, {"B200LK_STACK_DEPTH": "0"},
                                 {"B200LK_STACK_DEPTH": "1"}, {"B200LK_STACK_DEPTH": "3"}, {"B200LK_S4_SYM": "0"},
                                 {"B200LK_S4_MINB": "8"}, {"B200LK_S4_MINB": "12", "B200LK_STACK_DEPTH": "1"}, {},
                                 {"B200LK_S4_PREFETCH": "1"}, {"B200LK_S4_EVICT": "0"}, {"B200LK_S4_EVICT": "0", "B200LK_STACK_DEPTH": "1"},
                                 {"B200LK_S4_BULK": "1"},
                                 {"B200LK_S4_BULK": "1", "B200LK_S4_RING": "1"}, {"B200LK_S4_BULK": "1", "B200LK_S4_RING": "2"},
                                 {"B200LK_S4_BULK": "1", "B200LK_S4_RING": "4", "B200LK_S4_MINB": "6"},
                                 {"B200LK_S4_BUDGET": "37"}, {"B200LK_S4_BUDGET": "2", "B200LK_STACK_DEPTH": "1"}])
@pytest.mark.parametrize("safe", [False, True])
def test_dna_paths_agree_with_oracle(oracle_lib, monkeypatch, env, safe):
    """The 4-state traversal has three routes -- one pattern-stationary launch (default), the same with
    the shared-memory stack overflowing into global re-reads, and the level-batched kernels -- plus the
    general (non-SYM) kernel and the other register-budget instantiations of the SYM one, with the L1 prefetch of
    the next node's table rows, without the farthest-consumer replacement of stack entries, and with the node vectors leaving through cp.async.bulk (ring of 1 / 2 / 3 / 4
    shared-memory slots) instead of the threads' stores, and with the node list cut into many short launches (37 or 2
    nodes each: vectors crossing a cut are re-read from HBM); all of them must match the oracle, scale counts included
    (deep tree so that rescaling fires).
    The hooks are read when a context is created."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    tree, model, rates, props, aln = make_problem(300, 300, seed=5, mean_len=0.4, gaps=0.02)
    g, o = both(aln, model, rates, props, nslots=300, safe_numeric=safe)
    rg, ro = full_eval(g, tree, 3), full_eval(o, tree, 3)
    assert max(int(s.max()) for s in ro["scale"].values()) >= 1
    assert_parity(rg, ro, TOL)


def test_partial_retraversal_reads_resident_partials(oracle_lib):
    """After a full traversal only part of the list is recomputed (what a branch-length change does):
    children outside the list come from the resident slots."""
    tree, model, rates, props, aln = make_problem(40, 600, seed=12)
    g, o = both(aln, model, rates, props, nslots=40)
    full_eval(g, tree, 0), full_eval(o, tree, 0)
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    tail = tr[-7:]                       # the last seven nodes; their other children stay resident
    u, v = tail[0][0], tail[0][1]
    tree.set_length(u, v, 0.37)
    tr2 = W.traversal(tree, c, 0)[-7:]
    for e in (g, o):
        e.update_partials(tr2)
    L = tree.get_length(0, c)
    a = g.branch_lnl((LEAF, 0), (INTERNAL, tr2[-1][2]), L)
    b = o.branch_lnl((LEAF, 0), (INTERNAL, tr2[-1][2]), L)
    assert rel(a, b) <= TOL
    for e in tr2:
        assert np.array_equal(g.fetch_scale_num(e[2]), o.fetch_scale_num(e[2]))


@pytest.mark.parametrize("env", [{}, {"B200LK_NO_DMMA": "1"}])
@pytest.mark.parametrize("safe", [False, True])
def test_protein_tensor_path_deep_tree(oracle_lib, monkeypatch, env, safe):
    """20-state path: FP64 tensor-core tiles (default) and the generic kernels agree with the oracle on a
    tree deep enough for rescaling to fire; pattern count not a multiple of the 64-pattern tile."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    tree, model, rates, props, aln = make_problem(160, 333, nstates=20, seed=14, mean_len=0.5, gaps=0.02,
                                                  compress=False)
    g, o = both(aln, model, rates, props, nslots=160, safe_numeric=safe)
    rg, ro = full_eval(g, tree, 1), full_eval(o, tree, 1)
    assert max(int(s.max()) for s in ro["scale"].values()) >= 1
    assert_parity(rg, ro, TOL)


def test_protein_mixture_many_categories(oracle_lib):
    """C-series shape: 6 classes x 4 rates = 24 categories, i.e. six staged category groups per node."""
    tree, model, rates, props, aln = make_problem(12, 150, nstates=20, nmix=6, seed=15)
    for safe in (False, True):
        g, o = both(aln, model, rates, props, nslots=12, safe_numeric=safe)
        assert_parity(full_eval(g, tree, 0), full_eval(o, tree, 0), TOL)


@pytest.mark.parametrize("safe", [False, True])
def test_protein_mixture_deep_tree_rescales_across_category_groups(oracle_lib, safe):
    """Mixture launches split the category groups over CTAs (grid.z): the NORM_LH decision needs the maximum over
    ALL groups of a pattern.  Deep tree so that rescaling fires; scale counts must stay bit-exact."""
    tree, model, rates, props, aln = make_problem(150, 141, nstates=20, nmix=3, seed=19, mean_len=0.5, gaps=0.02,
                                                  compress=False)
    g, o = both(aln, model, rates, props, nslots=150, safe_numeric=safe)
    rg, ro = full_eval(g, tree, 2), full_eval(o, tree, 2)
    assert max(int(s.max()) for s in ro["scale"].values()) >= 1
    assert_parity(rg, ro, TOL)


def _newton_reference(eng, dad, node, guess, lo=1e-6, hi=10.0, max_steps=100):
    """minimizeNewton (utils/optimization.cpp:422-499) driven through build_theta/derv of an engine."""
    eng.build_theta(dad, node)
    evals = [0]

    def fd(t):
        evals[0] += 1
        df, ddf = eng.derv(t)
        return -df, -ddf
    rts = min(max(guess, lo), hi)
    f, df = fd(rts)
    if df >= 0.0 and abs(f) < lo:
        return rts, evals[0]
    if f < 0.0:
        xl, xh = rts, hi
    else:
        xh, xl = rts, lo
    for j in range(1, max_steps + 1):
        old = rts
        if df <= 0.0 or ((rts - xh) * df - f) * ((rts - xl) * df - f) >= 0.0:
            dx = 0.5 * (xh - xl)
            rts = xl + dx
            if xl == rts:
                return rts, evals[0]
        else:
            dx = f / df
            tmp = rts
            rts -= dx
            if tmp == rts:
                return rts, evals[0]
        if abs(dx) < lo or j == max_steps:
            return old, evals[0]
        f, df = fd(rts)
        if df > 0.0 and abs(f) < lo:
            return rts, evals[0]
        if f < 0.0:
            xl = rts
        elif f > 0.0:
            xh = rts
    raise AssertionError("not reached")


@pytest.mark.gpu
@pytest.mark.parametrize("nstates,safe", [(4, False), (4, True), (20, False)])
def test_device_newton_raphson_follows_the_reference_iteration(oracle_lib, nstates, safe):
    """b200lk_optimize_branch == the reference's minimizeNewton driven step by step through the oracle:
    same optimum, same number of derivative evaluations, from starts on both sides of the optimum."""
    tree, model, rates, props, aln = make_problem(10, 600 if nstates == 4 else 120, nstates=nstates, seed=17)
    gpu, cpu = both(aln, model, rates, props, nslots=tree.ntaxa, safe_numeric=safe)
    leaf = 0
    c = tree.adj[leaf][0]
    tr = W.traversal(tree, c, leaf)
    gpu.update_partials(tr)
    cpu.update_partials(tr)
    dad, node = (LEAF, leaf), (INTERNAL, tr[-1][2])
    for guess in (tree.get_length(leaf, c), 1e-6, 0.9, 5.0, 10.0):
        want, nev = _newton_reference(cpu, dad, node, guess)
        got, d2l, steps = gpu.optimize_branch(dad, node, guess)
        assert abs(got - want) <= 1e-9 * max(1.0, abs(want)), (guess, got, want)
        assert steps == nev, (guess, steps, nev)
        # theta stays valid: lnL from buffer at the optimum equals the oracle's
        assert abs(gpu.lnl_from_buffer(got) - cpu.lnl_from_buffer(want)) <= 1e-9 * abs(cpu.lnl_from_buffer(want))


@pytest.mark.gpu
@pytest.mark.parametrize("nstates,safe", [(4, False), (4, True), (20, False), (5, True)])
def test_ascertainment_bias_correction(oracle_lib, nstates, safe):
    """+ASC (Lewis): unobserved constant patterns appended after the observed ones; the log(1 - prob_const)
    correction reaches lnL, per-pattern lnL, df/ddf and lnL-from-buffer (oracle pinned against the reference's own
    kernel in tests/test_oracle_vs_ref.py::test_ascertainment_bias_correction_lewis)."""
    tree, model, rates, props, aln = make_problem(11, 333 if nstates == 4 else 77, nstates=nstates, seed=29)
    unobs = np.repeat(np.arange(nstates, dtype=np.uint8)[None, :], tree.ntaxa, axis=0)
    g, o = both(aln, model, rates, props, nslots=tree.ntaxa, safe_numeric=safe, unobserved=unobs)
    rg, ro = full_eval(g, tree, 0), full_eval(o, tree, 0)
    assert_parity(rg, ro, TOL)
    for t in (0.01, 0.3):
        a, b = g.lnl_from_buffer(t), o.lnl_from_buffer(t)
        assert abs(a - b) <= TOL * abs(b)
        da, db = g.derv(t), o.derv(t)
        assert all(abs(x - y) <= TOL * max(1.0, abs(y)) for x, y in zip(da, db))


def test_ufboot_rell_from_resident_pattern_lh(oracle_lib):
    """UFBoot RELL (IQTree::saveCurrentTree, tree/iqtree.cpp:3606-3616): rell[b] = dotProduct(pattern_lh, boot_samples[b])
    in the reference's BootValType = float.  Restated here with numpy on the oracle's per-pattern lnL; the float
    products are identical, the sum is accumulated in double on the device, so agreement is bounded by the float
    rounding of the reference's own accumulation: |d| <= 1e-6 * sum |terms|."""
    tree, model, rates, props, aln = make_problem(10, 3000, seed=33, compress=False)
    g, o = both(aln, model, rates, props, nslots=tree.ntaxa)
    rg, ro = full_eval(g, tree, 0), full_eval(o, tree, 0)
    rng = np.random.default_rng(5)
    nsite = int(aln.ptn_freq.sum())
    samples = rng.multinomial(nsite, aln.ptn_freq / nsite, size=200).astype(np.float32)  # Alignment::createBootstrapAlignment
    g.upload_boot_samples(samples)
    got = g.rell()
    plh32 = ro["pattern_lh"].astype(np.float32)
    want = (plh32[None, :] * samples).astype(np.float64).sum(axis=1)
    bound = 1e-6 * (np.abs(plh32[None, :] * samples)).astype(np.float64).sum(axis=1)
    assert np.all(np.abs(got - want) <= bound)
    # the original alignment as a "resample" reproduces lnL to float precision
    g.upload_boot_samples(aln.ptn_freq[None, :].astype(np.float32))
    assert abs(g.rell()[0] - rg["lnl"]) <= 1e-6 * abs(rg["lnl"])
    # RELL follows the tree: after lnL-from-buffer at another length the resident per-pattern lnL changed
    g.build_theta((LEAF, 0), (INTERNAL, rg["top"]))
    other = g.lnl_from_buffer(0.77)
    assert abs(g.rell()[0] - other) <= 1e-6 * abs(other)


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("nstates,ntaxa,mean_len", [(4, 30, 0.1), (4, 150, 0.4), (20, 24, 0.1), (20, 90, 0.5)])
def test_batched_launches_over_partitions_match_per_partition_results(oracle_lib, nstates, ntaxa, mean_len, safe):
    """b200lk_batch_*: genes of a supermatrix on one topology, own alignments (different pattern counts), models and
    branch lengths, evaluated by batched launches -- against the oracle gene by gene
    (the loop of tree/phylosupertree.cpp:701-725)."""
    import copy
    from iqtree2_b200.capi import Batch
    ngenes = 5
    rng = np.random.default_rng(77)
    base, _, rates, props, _ = make_problem(ntaxa, 10, nstates=nstates, seed=31, mean_len=mean_len)
    trees, engines, oracles = [], [], []
    for gene in range(ngenes):
        t = copy.deepcopy(base)
        for (u, v) in t.branches():
            t.set_length(u, v, base.get_length(u, v) * (0.5 + gene * 0.3) * rng.uniform(0.8, 1.25))
        model = W.gtr_model(rates=tuple(rng.uniform(0.5, 3.0, 6))) if nstates == 4 else \
            W.random_reversible_model(nstates, 500 + gene)
        aln = W.simulate_alignment(t, model, rates, props, 120 + 173 * gene, 900 + gene)
        g, o = both(aln, model, rates, props, nslots=ntaxa, safe_numeric=safe)
        trees.append(t)
        engines.append(g)
        oracles.append(o)
    leaf = 3
    c = base.adj[leaf][0]
    tr = W.traversal(base, c, leaf)
    top = tr[-1][2]
    child_len = np.array([[t.get_length(e[0], ch_node) for e, ch_node in _children_of(base, tr)] for t in trees])
    blen = np.array([t.get_length(leaf, c) for t in trees])
    batch = Batch(engines)
    lnl, df, ddf = batch.evaluate(tr, child_len, (LEAF, leaf), (INTERNAL, top), blen)
    pattern_lh = [e.fetch_pattern_lh() for e in engines]
    blen2 = blen * 1.7
    df2, ddf2 = batch.derv(blen2)
    lnl2 = batch.lnl_from_buffer(blen2)
    if ntaxa >= 90:
        assert max(int(engines[-1].fetch_scale_num(s).max()) for (_, _, s, _) in tr) >= 1
    for gene in range(ngenes):
        want = full_eval(oracles[gene], trees[gene], leaf)
        assert rel(lnl[gene], want["lnl"]) <= TOL and rel(df[gene], want["df"]) <= TOL and rel(ddf[gene], want["ddf"]) <= TOL
        for (_, _, s, _) in tr:
            assert np.array_equal(engines[gene].fetch_scale_num(s), want["scale"][s])
        assert rel(pattern_lh[gene], want["pattern_lh"]) <= TOL
        wdf, wddf = oracles[gene].derv(blen2[gene])
        assert rel(df2[gene], wdf) <= TOL and rel(ddf2[gene], wddf) <= TOL
        assert rel(lnl2[gene], oracles[gene].lnl_from_buffer(blen2[gene])) <= TOL
        # the per-context calls still work on a context that belongs to a batch
        assert rel(engines[gene].lnl_from_buffer(blen2[gene]), lnl2[gene]) <= 1e-12


def _children_of(tree, tr):
    """(entry, neighbour node) for every child of a traversal list, in child order."""
    out = []
    for e in tr:
        node, dad = e[0], e[1]
        kids = [w for w in tree.adj[node] if w != dad]
        assert len(kids) == len(e[3])
        out += [(e, w) for w in kids]
    return out


def test_batch_refuses_what_it_cannot_express(oracle_lib):
    from iqtree2_b200.capi import B200LKError, Batch
    tree, model, rates, props, aln = make_problem(8, 100, nstates=5, ncat=2, seed=3)
    g = engine_for(aln, model, rates, props, nslots=8)
    with pytest.raises(B200LKError, match="-5"):
        Batch([g])
    tree, model, rates, props, aln = make_problem(8, 100, seed=3)
    a = engine_for(aln, model, rates, props, nslots=8)
    b = engine_for(aln, model, rates, props, nslots=8, safe_numeric=True)
    with pytest.raises(B200LKError, match="-5"):
        Batch([a, b])


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("nstates,ncat,nmix,ntaxa,nsite", [
    (9, 2, 1, 8, 150), (12, 3, 2, 9, 130), (16, 1, 1, 6, 100), (17, 2, 1, 7, 90), (24, 2, 1, 7, 90), (32, 2, 1, 6, 70),
    (33, 1, 1, 6, 70), (40, 2, 1, 6, 60), (48, 1, 1, 5, 60), (56, 1, 2, 5, 50), (61, 4, 1, 8, 130), (64, 1, 1, 5, 45),
])
def test_generic_tensor_path_all_tile_counts(oracle_lib, safe, nstates, ncat, nmix, ntaxa, nsite):
    """densen.cu: 9 <= nstates <= 64 (codon 61, padded to 8-state tiles) on the FP64 tensor cores."""
    tree, model, rates, props, aln = make_problem(ntaxa, nsite, nstates=nstates, ncat=ncat, nmix=nmix, seed=41,
                                                  gaps=0.08)
    g, o = both(aln, model, rates, props, nslots=ntaxa, safe_numeric=safe)
    for leaf in (0, ntaxa - 1):
        assert_parity(full_eval(g, tree, leaf), full_eval(o, tree, leaf), TOL)


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("nstates", [13, 61])
def test_generic_tensor_path_deep_tree_rescales(oracle_lib, nstates, safe):
    ntaxa = 140 if nstates == 13 else 70
    tree, model, rates, props, aln = make_problem(ntaxa, 300 if nstates == 13 else 150, nstates=nstates, ncat=2, seed=9,
                                                  mean_len=0.6)
    g, o = both(aln, model, rates, props, nslots=ntaxa, safe_numeric=safe)
    rg, ro = full_eval(g, tree, 0), full_eval(o, tree, 0)
    assert max(int(s.max()) for s in ro["scale"].values()) >= 1
    assert_parity(rg, ro, TOL)


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("nstates,k", [(4, 4), (4, 2), (20, 3), (12, 2), (6, 2)])
def test_fused_mixture_rate_categories(oracle_lib, nstates, k, safe):
    """fused_mix_rate: category c = mixture class c at rate category c, on every kernel family."""
    tree, model, rates, props, aln = make_problem(11, 260, nstates=nstates, ncat=k, nmix=k, seed=14, gaps=0.04)
    g, o = both(aln, model, rates, props, nslots=11, safe_numeric=safe, fused=True)
    assert g.ncat_mix == k
    assert_parity(full_eval(g, tree, 2), full_eval(o, tree, 2), TOL)


def _alternating_polytomy_newick(ntaxa):
    """Caterpillar whose internal nodes alternate between two and three children: every run of bifurcating nodes is
    one node long, so the 4-state traversal is cut into as many launches as a tree can force."""
    text, used, tri = "(T0:0.1,T1:0.12)", 2, True
    while used < ntaxa - 2:
        if tri and used + 2 <= ntaxa - 2:
            text = "(%s:0.05,T%d:0.1,T%d:0.07)" % (text, used, used + 1)
            used += 2
        else:
            text = "(%s:0.04,T%d:0.09)" % (text, used)
            used += 1
        tri = not tri
    return "(%s:0.03,T%d:0.1,T%d:0.2);" % (text, used, used + 1)


@pytest.mark.parametrize("safe", [False, True])
def test_polytomy_heavy_tree_tip_state_staging(oracle_lib, safe):
    """every launch of the 4-state traversal used to pad its tip-state row sequence to a whole chunk, and a
    tree of ~550+ taxa with alternating bi-/trifurcations overflowed the staging buffer.  The padding is now shared
    across launches and the capacity is checked."""
    ntaxa = 640
    names = ["T%d" % i for i in range(ntaxa)]
    t = W.Tree.from_newick(_alternating_polytomy_newick(ntaxa), names)
    model = W.gtr_model()
    rates, props = W.discrete_gamma(0.7)
    aln = W.simulate_alignment(W.random_tree(ntaxa, 4, mean_len=0.02), model, rates, props, 97, 11, compress=True)
    aln.names = names
    g, o = both(aln, model, rates, props, nslots=ntaxa, safe_numeric=safe)
    rg, ro = full_eval(g, t, 0), full_eval(o, t, 0)
    assert sum(1 for e in ro["traversal"] if len(e[3]) > 2) > 150
    assert_parity(rg, ro, TOL)
    # a second evaluation reuses the staging ring slices behind the first one's
    assert_parity(full_eval(g, t, ntaxa - 1), full_eval(o, t, ntaxa - 1), TOL)


@pytest.mark.parametrize("nstates", [4, 20])
def test_batched_evaluation_repairs_non_finite_lnl_like_the_per_context_path(oracle_lib, nstates):
    """one gene with a NaN pattern must not poison a batched supermatrix evaluation: the reference replaces
    non-finite per-pattern values by 4 * LOG_SCALING_THRESHOLD and re-sums (tree/phylokernelnew.h:3077-3094), the
    per-context path does the same (finish_lnl), and so must the batch."""
    from iqtree2_b200.capi import WARN_LNL_NONFINITE, Batch
    ntaxa, ngenes = 9, 3
    base, _, rates, props, _ = make_problem(ntaxa, 10, nstates=nstates, seed=31)
    engines, oracles = [], []
    for gene in range(ngenes):
        model = W.gtr_model() if nstates == 4 else W.random_reversible_model(nstates, 500 + gene)
        aln = W.simulate_alignment(base, model, rates, props, 150 + 40 * gene, 900 + gene)
        g, o = both(aln, model, rates, props, nslots=ntaxa)
        if gene == 1:  # poison pattern 5 of gene 1: lh = |x| + NaN
            inv = np.zeros(aln.nptn)
            inv[5] = np.nan
            g.set_ptn_invar(inv)
            o.set_ptn_invar(inv)
        engines.append(g)
        oracles.append(o)
    leaf = 2
    c = base.adj[leaf][0]
    tr = W.traversal(base, c, leaf)
    top = tr[-1][2]
    L = base.get_length(leaf, c)
    child_len = np.array([[base.get_length(e[0], ch_node) for e, ch_node in _children_of(base, tr)]] * ngenes)
    blen = np.full(ngenes, L)
    lnl, _, _ = Batch(engines).evaluate(tr, child_len, (LEAF, leaf), (INTERNAL, top), blen)
    for gene in range(ngenes):
        oracles[gene].update_partials(tr)
        want = oracles[gene].branch_lnl((LEAF, leaf), (INTERNAL, top), L)
        alone = engines[gene].evaluate(tr, (LEAF, leaf), (INTERNAL, top), L)[0]  # unbatched path, same context
        assert np.isfinite(want) and np.isfinite(lnl[gene])
        assert rel(lnl[gene], want) <= TOL and rel(alone, want) <= TOL
        if gene == 1:
            assert engines[gene].last_status & WARN_LNL_NONFINITE
            assert engines[gene].fetch_pattern_lh()[5] == 4 * -177.4456782233459932741


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("ncat,nmix,p_invar", [(4, 1, 0.0), (2, 2, 0.0), (3, 1, 0.15), (1, 1, 0.0)])
def test_fused_evaluation_tail_equals_the_separate_kernels(oracle_lib, safe, ncat, nmix, p_invar):
    """b200lk_evaluate on the 4-state layout runs lnL + theta + df/ddf as ONE launch (eval_tail_kernel).  Its per-pattern
    outputs must be bit-identical to the separate lnl / theta kernels', its sums equal to theirs and to the oracle's."""
    tree, model, rates, props, aln = make_problem(40, 3000, ncat=ncat, nmix=nmix, seed=51, gaps=0.05, mean_len=0.4)
    g, o = both(aln, model, rates, props, nslots=40, safe_numeric=safe, p_invar=p_invar)
    internal = next(v for v in range(tree.ntaxa, tree.nnodes) if all(not tree.is_leaf(u) for u in tree.adj[v][:2]))
    for dad_node, node in ((0, tree.adj[0][0]), (internal, tree.adj[internal][0])):
        tr = W.traversal(tree, node, dad_node) + ([] if tree.is_leaf(dad_node) else W.traversal(tree, dad_node, node))
        slot = {(e[0], e[1]): e[2] for e in tr}
        if tree.is_leaf(dad_node):
            dad, nd = (LEAF, dad_node), (INTERNAL, slot[(node, dad_node)])
        else:
            tr = _two_sided(tree, node, dad_node)
            slot = {(e[0], e[1]): e[2] for e in tr}
            dad, nd = (INTERNAL, slot[(dad_node, node)]), (INTERNAL, slot[(node, dad_node)])
        L = tree.get_length(dad_node, node)
        fused = g.evaluate(tr, dad, nd, L)
        f_plh, f_cat = g.fetch_pattern_lh(), g.fetch_pattern_lh_cat()
        f_theta, f_bs = g.fetch_theta()
        lnl = g.branch_lnl(dad, nd, L)
        s_plh, s_cat = g.fetch_pattern_lh(), g.fetch_pattern_lh_cat()
        g.build_theta(dad, nd)
        s_theta, s_bs = g.fetch_theta()
        df, ddf = g.derv(L)
        assert np.array_equal(f_plh, s_plh) and np.array_equal(f_cat, s_cat)
        assert np.array_equal(f_theta, s_theta) and np.array_equal(f_bs, s_bs)
        assert rel(fused, (lnl, df, ddf)) <= 1e-12
        assert rel(fused, o.evaluate(tr, dad, nd, L)) <= TOL
        assert g.evaluate(tr, dad, nd, L) == fused  # deterministic last-block reduction


def _two_sided(tree, node, dad):
    """post-order lists of both directions of branch (dad, node) with distinct slots"""
    a = W.traversal(tree, node, dad)
    used = {e[2] for e in a}
    free = iter(s for s in range(tree.ntaxa) if s not in used)
    remap = {}
    b = []
    for (v, u, s, ch) in W.traversal(tree, dad, node):
        remap[s] = next(free)
        b.append((v, u, remap[s], [(k, remap[i] if k == INTERNAL else i, l) for k, i, l in ch]))
    return a + b


def test_status_bit_marks_the_documented_derivative_deviation(oracle_lib):
    """INTEGRATION.md 'Deviations': with odd nstates and an even number of categories the reference counts the last block
    entry twice in df/ddf; the library returns the true sums and says so with B200LK_INFO_DERV_NOT_AS_REFERENCE."""
    from iqtree2_b200.capi import INFO_DERV_NOT_AS_REFERENCE
    for nstates, ncat, expect in ((5, 2, True), (5, 3, False), (4, 4, False), (20, 4, False)):
        tree, model, rates, props, aln = make_problem(6, 60, nstates=nstates, ncat=ncat, seed=3)
        g = engine_for(aln, model, rates, props, nslots=6)
        c = tree.adj[0][0]
        tr = W.traversal(tree, c, 0)
        g.evaluate(tr, (LEAF, 0), (INTERNAL, tr[-1][2]), 0.1)
        assert bool(g.last_status & INFO_DERV_NOT_AS_REFERENCE) == expect
        g.derv(0.2)
        assert bool(g.last_status & INFO_DERV_NOT_AS_REFERENCE) == expect
        g.branch_lnl((LEAF, 0), (INTERNAL, tr[-1][2]), 0.1)
        assert g.last_status == 0


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("ncat,nmix", [(4, 1), (2, 2)])  # (2, 2): two model classes -> the general (non-SYM) kernel
def test_tip_tip_state_pair_tables_change_no_bit(oracle_lib, safe, ncat, nmix):
    """4-state traversal: with clearly more patterns than state pairs a tip-tip node is looked up in a per-node state-pair
    table (s4_tt_tables_kernel).  Patterns are independent, so the first 1000 patterns evaluated inside a 3000-pattern
    context (tables on) and alone (tables off: 1000 < 4 * 19^2) must give bit-identical vectors and scale counts."""
    tree, model, rates, props, aln = make_problem(30, 3000, ncat=ncat, nmix=nmix, seed=61, gaps=0.08, compress=False)
    sub = W.Alignment(aln.nstates, aln.names, np.ascontiguousarray(aln.states[:, :1000]), np.ascontiguousarray(aln.ptn_freq[:1000]))
    big, o = both(aln, model, rates, props, nslots=30, safe_numeric=safe)
    small = engine_for(sub, model, rates, props, nslots=30, safe_numeric=safe)
    rb, rs = full_eval(big, tree, 0), full_eval(small, tree, 0)
    assert sum(1 for e in rb["traversal"] if all(ch[0] == LEAF for ch in e[3])) >= 5  # there are tip-tip nodes
    for (_, _, s, _) in rb["traversal"]:
        assert np.array_equal(big.fetch_partial(s)[:1000], small.fetch_partial(s))
        assert np.array_equal(big.fetch_scale_num(s)[:1000], small.fetch_scale_num(s))
    assert np.array_equal(rb["pattern_lh"][:1000], rs["pattern_lh"])
    assert_parity(rb, full_eval(o, tree, 0), TOL)


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("ncat", [1, 4])
@pytest.mark.parametrize("nptn", [1000, 1017, 1056])  # whole tiles / a ragged last warp / a whole last warp + an empty one
def test_bulk_store_changes_no_bit(oracle_lib, monkeypatch, safe, ncat, nptn):
    """B200LK_S4_BULK=1: the node vectors of the SYM traversal leave through cp.async.bulk from a shared-memory image of
    the warp's tile (s4_traverse_kernel<.., BULK>); the arithmetic is untouched, so every vector, scale count and pattern
    likelihood must be bit-identical to the default build that stores from the threads (with and without the
    farthest-consumer replacement of stack entries and the L1 prefetch, which change the order of memory operations only) -- also with a ring of one slot (every non-adjacent consumer re-reads what the copy engine wrote) and
    when the pattern count leaves a ragged or an empty warp."""
    tree, model, rates, props, aln = make_problem(120, nptn, ncat=ncat, seed=83, mean_len=0.3, gaps=0.05, compress=False)
    res = {}
    for tag, env in (("stg", {"B200LK_S4_EVICT": "0"}), ("evict", {}), ("prefetch", {"B200LK_S4_PREFETCH": "1"}), ("bulk", {"B200LK_S4_BULK": "1"}),
                     ("ring1", {"B200LK_S4_BULK": "1", "B200LK_S4_RING": "1"})):
        with monkeypatch.context() as m:
            for k, v in env.items():
                m.setenv(k, v)
            g = engine_for(aln, model, rates, props, nslots=120, safe_numeric=safe)
        r = full_eval(g, tree, 0)
        res[tag] = (r, {s: (g.fetch_partial(s), g.fetch_scale_num(s)) for (_, _, s, _) in r["traversal"]})
    (r0, v0) = res["stg"]
    for tag in ("evict", "prefetch", "bulk", "ring1"):
        r, v = res[tag]
        for s in v0:
            assert np.array_equal(v[s][0], v0[s][0]), (tag, s)
            assert np.array_equal(v[s][1], v0[s][1]), (tag, s)
        assert np.array_equal(r["pattern_lh"], r0["pattern_lh"])
        assert r["lnl"] == r0["lnl"] and r["df"] == r0["df"] and r["ddf"] == r0["ddf"]


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("ncat", [1, 4])
def test_virtual_tip_tip_vectors_change_no_bit(oracle_lib, monkeypatch, safe, ncat):
    """A tip-tip vector that the traversal's next consumer takes from registers or the stack is not written by
    s4_traverse_kernel: the slot stays virtual until something else needs it (s4_realize gathers it from the state-pair
    table).  Every slot, scale count and result must be bit-identical to the build that writes everything
    (B200LK_S4_VIRTUAL=0) -- read back right after the traversal, after a second traversal with other branch lengths,
    and after a partial re-traversal whose children outside the list are virtual slots."""
    monkeypatch.setenv("B200LK_S4_TT", "2")
    tree, model, rates, props, aln = make_problem(60, 1500, ncat=ncat, seed=29, mean_len=0.25, gaps=0.05, compress=False)
    out = {}
    for tag, env in (("real", {"B200LK_S4_VIRTUAL": "0"}), ("virtual", {})):
        with monkeypatch.context() as m:
            for k, v in env.items():
                m.setenv(k, v)
            g = engine_for(aln, model, rates, props, nslots=60, safe_numeric=safe)
        t = W.Tree.from_newick(tree.to_newick(), aln.names)  # a private copy: the lengths change below
        c = t.adj[0][0]
        snaps = []
        r = full_eval(g, t, 0)
        tr = r["traversal"]
        assert sum(1 for e in tr if all(ch[0] == LEAF for ch in e[3])) >= 8
        snaps.append((r["lnl"], r["df"], r["ddf"], r["pattern_lh"], {s: (g.fetch_partial(s), g.fetch_scale_num(s)) for (_, _, s, _) in tr}))
        # second traversal, other lengths: every virtual slot is overwritten, nothing is gathered
        for (u, v) in list(t.branches()):
            t.set_length(u, v, 0.7 * t.get_length(u, v) + 0.01)
        g.update_partials(W.traversal(t, c, 0))
        # partial re-traversal: the last five nodes; their other children (virtual tip-tip slots among them) stay resident
        tail = W.traversal(t, c, 0)[-5:]
        t.set_length(tail[0][0], tail[0][1], 0.33)
        tr2 = W.traversal(t, c, 0)[-5:]
        g.update_partials(tr2)
        lnl = g.branch_lnl((LEAF, 0), (INTERNAL, tr2[-1][2]), t.get_length(0, c))
        snaps.append((lnl, g.fetch_pattern_lh(), {s: (g.fetch_partial(s), g.fetch_scale_num(s)) for (_, _, s, _) in tr}))
        out[tag] = snaps
    a, b = out["real"], out["virtual"]
    assert a[0][0] == b[0][0] and a[0][1] == b[0][1] and a[0][2] == b[0][2]
    assert np.array_equal(a[0][3], b[0][3])
    for s in a[0][4]:
        assert np.array_equal(a[0][4][s][0], b[0][4][s][0]) and np.array_equal(a[0][4][s][1], b[0][4][s][1]), s
    assert a[1][0] == b[1][0] and np.array_equal(a[1][1], b[1][1])
    for s in a[1][2]:
        assert np.array_equal(a[1][2][s][0], b[1][2][s][0]) and np.array_equal(a[1][2][s][1], b[1][2][s][1]), s


def test_virtual_slot_read_by_a_branch_kernel(oracle_lib, monkeypatch):
    """lnL and Newton-Raphson on the branch ABOVE a tip-tip node right after the traversal that left it virtual: the
    branch kernels gather the slot first (fill_args -> s4_realize) and agree with the oracle."""
    monkeypatch.setenv("B200LK_S4_TT", "2")
    names = list("ABCDEF")
    t = W.Tree.from_newick("((A:0.1,B:0.2):0.15,(C:0.1,D:0.3):0.05,(E:0.2,F:0.1):0.3);", names)
    model = W.gtr_model()
    rates, props = W.discrete_gamma(0.7)
    aln = W.simulate_alignment(W.random_tree(6, 4), model, rates, props, 700, 11, compress=False)
    aln.names = names
    g, o = both(aln, model, rates, props, nslots=12)
    root = [u for u in range(t.nnodes) if not t.is_leaf(u) and sum(1 for w in t.adj[u] if not t.is_leaf(w)) == 3][0]
    kids = t.adj[root]
    # slots: cherry k -> slot k (seen from the root), root seen from cherry 0 -> slot 5 (children: cherries 1 and 2)
    cherries = []
    for k, v in enumerate(kids):
        leaves = [w for w in t.adj[v] if w != root]
        cherries.append((v, root, k, [(LEAF, w, t.get_length(v, w)) for w in leaves]))
    top = (root, kids[0], 5, [(INTERNAL, 1, t.get_length(root, kids[1])), (INTERNAL, 2, t.get_length(root, kids[2]))])
    top1 = (root, kids[1], 6, [(INTERNAL, 0, t.get_length(root, kids[0])), (INTERNAL, 2, t.get_length(root, kids[2]))])
    tr = cherries + [top, top1]
    L = t.get_length(root, kids[0])
    for e in (g, o):
        e.update_partials(tr)
    # cherry 2 (slot 2) has two readers in the list: stored.  Cherries 0 and 1 have one reader each, which takes them from
    # the stack: virtual.  The branch kernels below read slot 0.
    a = g.branch_lnl((INTERNAL, 5), (INTERNAL, 0), L)
    b = o.branch_lnl((INTERNAL, 5), (INTERNAL, 0), L)
    assert rel(a, b) <= TOL
    a = g.branch_lnl((INTERNAL, 0), (INTERNAL, 5), L)
    assert rel(a, b) <= TOL
    L1 = t.get_length(root, kids[1])
    assert rel(g.branch_lnl((INTERNAL, 6), (INTERNAL, 1), L1), o.branch_lnl((INTERNAL, 6), (INTERNAL, 1), L1)) <= TOL
    for s in (0, 1, 2, 5, 6):
        assert rel(g.fetch_partial(s), o.fetch_partial(s)) <= TOL
        assert np.array_equal(g.fetch_scale_num(s), o.fetch_scale_num(s))


@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("ncat,nmix", [(4, 1), (2, 3)])  # (2, 3): six categories -> the split (mixture) launch and its fix-up kernel
def test_virtual_tip_tip_vectors_20_state(oracle_lib, monkeypatch, safe, ncat, nmix):
    """20-state path: a tip-tip node whose only reader is a bifurcating node of the tensor kernel is not gathered into its
    slot; the reader takes its operand rows from the node's state-pair table by per-pattern pair index
    (dense20_rp_kernel, Lvirt / Rvirt).  Same bits as the build that gathers every node (B200LK_S4_VIRTUAL=0), parity with
    the oracle, and the slots read back afterwards (s4_realize -> dense_realize) hold the same vectors."""
    monkeypatch.setenv("B200LK_S4_TT", "2")
    tree, model, rates, props, aln = make_problem(24, 700, nstates=20, ncat=ncat, nmix=nmix, seed=37, mean_len=0.3, gaps=0.04,
                                                  compress=False)
    out = {}
    for tag, env in (("real", {"B200LK_S4_VIRTUAL": "0"}), ("virtual", {})):
        with monkeypatch.context() as m:
            for k, v in env.items():
                m.setenv(k, v)
            g = engine_for(aln, model, rates, props, nslots=24, safe_numeric=safe)
        r = full_eval(g, tree, 0)
        tr = r["traversal"]
        assert sum(1 for e in tr if all(ch[0] == LEAF for ch in e[3])) >= 4
        vecs = {s: (g.fetch_partial(s), g.fetch_scale_num(s)) for (_, _, s, _) in tr}
        # a second evaluation from the other end, into other slots, reads the (realized) vectors of the first where it can
        r2 = full_eval(g, tree, 5, slot_base=0)
        out[tag] = (r, vecs, r2)
    (ra, va, ra2), (rb, vb, rb2) = out["real"], out["virtual"]
    assert ra["lnl"] == rb["lnl"] and ra["df"] == rb["df"] and ra["ddf"] == rb["ddf"]
    assert np.array_equal(ra["pattern_lh"], rb["pattern_lh"])
    for s in va:
        assert np.array_equal(va[s][0], vb[s][0]) and np.array_equal(va[s][1], vb[s][1]), s
    assert ra2["lnl"] == rb2["lnl"] and np.array_equal(ra2["pattern_lh"], rb2["pattern_lh"])
    oracle.set_ref_quirks(False)
    o = oracle.oracle_engine(aln, model, rates, props, nslots=24, safe_numeric=safe)
    assert_parity(rb, full_eval(o, tree, 0), TOL)
