"""Site-specific models (SITE_MODEL = true kernel variants, ModelSet / PMSF): the oracle's
restatement against the reference's own SITE_MODEL kernels (oracle/_ref), and the CUDA path against the oracle."""
import numpy as np
import pytest

import oracle
from iqtree2_b200 import workload as W
from tests.common import assert_parity, full_eval, make_problem, rel, site_engine, site_models_for
from tests.test_oracle_vs_ref import leaf_sets


@pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("kernel", [oracle.LK_AVX_FMA, oracle.LK_SSE2])
@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("nstates,ncat,ntaxa,nsite,mean_len", [(20, 4, 9, 61, 0.1), (4, 3, 12, 150, 0.1), (6, 2, 8, 50, 0.1),
                                                                (20, 2, 120, 30, 0.5)])
def test_oracle_site_models_match_reference(oracle_lib, kernel, safe, nstates, ncat, ntaxa, nsite, mean_len):
    tree, _, rates, props, aln = make_problem(ntaxa, nsite, nstates=nstates, ncat=ncat, seed=61, gaps=0.06, mean_len=mean_len)
    models = site_models_for(aln, nstates, 62)
    dummy = W.random_reversible_model(nstates, 1)
    ref = oracle.RefRun(tree, aln, dummy, rates, props, kernel=kernel, safe=safe, site_models=models)
    eng = site_engine(aln, models, rates, props, nslots=ntaxa, safe_numeric=ref.safe_numeric, lib_path=oracle.ORACLE_LIB,
                      prefix="orc_")
    leaf = 2
    lnl = ref.lnl_at_leaf(leaf)
    o = full_eval(eng, tree, leaf)
    assert rel(lnl, o["lnl"]) <= 1e-11
    assert rel(ref.pattern_lh(), o["pattern_lh"]) <= 1e-11
    cat_r, cat_o = ref.pattern_lh_cat(), o["pattern_lh_cat"]
    denom = np.maximum(np.abs(cat_o).max(axis=1, keepdims=True), 1e-300)
    assert float(np.max(np.abs(cat_r - cat_o) / denom)) <= 1e-9
    df, ddf = ref.derv()
    assert rel(df, o["df"]) <= 1e-9 and rel(ddf, o["ddf"]) <= 1e-9
    assert rel(ref.from_buffer(), o["lnl_buf"]) <= 1e-11
    nz = 0
    for slot, ls in leaf_sets(tree, o["traversal"]).items():
        part, sc = ref.fetch_branch(ls, leaf)
        assert np.array_equal(sc, o["scale"][slot]), "scale_num differs from the reference at slot %d" % slot
        nz += int(sc.max() > 0)
        got = eng.fetch_partial(slot)
        d = np.maximum(np.abs(part).reshape(part.shape[0], -1).max(axis=1), 1e-300)[:, None, None]
        assert float(np.max(np.abs(got - part) / d)) <= 1e-9
    if ntaxa >= 100:
        assert nz > 0, "the deep tree must make rescaling fire"
    L = tree.get_length(leaf, tree.adj[leaf][0])
    for t in (0.4 * L, 2.5 * L):
        ref.set_current_length(t)
        assert rel(ref.derv(again=True), eng.derv(t)) <= 1e-9
        assert rel(ref.from_buffer(), eng.lnl_from_buffer(t)) <= 1e-11
    ref.close()


def test_site_models_reduce_to_the_homogeneous_model(oracle_lib):
    """Every pattern given the SAME eigen-system: the site-specific path must reproduce the ordinary one."""
    tree, model, rates, props, aln = make_problem(10, 200, nstates=20, seed=8, gaps=0.05)
    models = tuple(np.repeat(a, aln.nptn, axis=0) for a in (model.eval, model.evec, model.inv_evec))
    site = site_engine(aln, models, rates, props, nslots=10, lib_path=oracle.ORACLE_LIB, prefix="orc_")
    plain = oracle.oracle_engine(aln, model, rates, props, nslots=10)
    a, b = full_eval(site, tree, 0), full_eval(plain, tree, 0)
    for k in ("lnl", "df", "ddf", "lnl_buf"):
        assert rel(a[k], b[k]) <= 1e-11, k
    assert rel(a["pattern_lh"], b["pattern_lh"]) <= 1e-11


@pytest.mark.gpu
@pytest.mark.parametrize("safe", [False, True])
@pytest.mark.parametrize("nstates,ncat,ntaxa,nsite,mean_len", [(20, 4, 10, 90, 0.1), (4, 4, 12, 200, 0.1), (4, 1, 6, 64, 0.1), (7, 3, 8, 70, 0.1),
                                                                (32, 2, 6, 40, 0.1), (20, 2, 130, 70, 0.5), (20, 9, 7, 40, 0.1)])
def test_site_models_cuda_matches_oracle(oracle_lib, safe, nstates, ncat, ntaxa, nsite, mean_len):
    oracle.set_ref_quirks(False)
    tree, _, rates, props, aln = make_problem(ntaxa, nsite, nstates=nstates, ncat=ncat, seed=71, gaps=0.06, mean_len=mean_len)
    models = site_models_for(aln, nstates, 72)
    g = site_engine(aln, models, rates, props, nslots=ntaxa, safe_numeric=safe)
    o = site_engine(aln, models, rates, props, nslots=ntaxa, safe_numeric=safe, lib_path=oracle.ORACLE_LIB, prefix="orc_")
    for leaf in (0, ntaxa - 1):
        rg, ro = full_eval(g, tree, leaf), full_eval(o, tree, leaf)
        assert_parity(rg, ro, 1e-9)
    if ntaxa >= 100:
        assert max(int(s.max()) for s in ro["scale"].values()) >= 1


@pytest.mark.gpu
def test_site_context_refuses_what_it_cannot_express(oracle_lib):
    from iqtree2_b200.capi import B200LKError, Engine
    with pytest.raises(B200LKError, match="-5"):
        Engine(ntaxa=4, nstates=20, ncat=2, nmix=2, nptn=10, nslots=4, n_state_codes=24, site_specific=True)
    tree, model, rates, props, aln = make_problem(6, 50, nstates=20, seed=3)
    g = site_engine(aln, site_models_for(aln, 20, 4), rates, props, nslots=6)
    with pytest.raises(B200LKError, match="-3"):
        g.set_model(model.eval, model.evec, model.inv_evec, rates, props, model.mixw, W.tip_partial_lh(model))
