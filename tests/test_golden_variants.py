"""Golden vectors of the variant paths (heterotachy, Holder's +ASC, robust scoring) computed by the REFERENCE's own
kernels (tests/golden/make_golden_variants.py, run where /root/reference exists), checked against the CPU oracle
(-m "not gpu": pins the oracle where the reference is absent) and the CUDA path through the C ABI (-m gpu).
Bar: 1e-9 relative (north_star)."""
import glob
import os

import numpy as np
import pytest

from iqtree2_b200 import workload as W
from iqtree2_b200.capi import INTERNAL, LEAF, engine_for
from tests.common import rel
from tests.test_mixlen_holder import holder_unobserved, mix_child_len

HERE = os.path.dirname(os.path.abspath(__file__))
FILES = sorted(glob.glob(os.path.join(HERE, "golden", "var_*.npz")))
TOL = 1e-9


def load(path):
    g = np.load(path, allow_pickle=False)
    if str(g["kind"]) == "nonrev":
        return g, None, None, None
    names = [str(x) for x in g["names"]]
    tree = W.Tree.from_newick(str(g["newick"]), names)
    model = W.EigenModel(nstates=int(g["nstates"]), eval=g["eval"], evec=g["evec"], inv_evec=g["inv_evec"],
                         freqs=g["freqs"], mixw=g["mixw"], Q=np.zeros(0))
    aln = W.Alignment(int(g["nstates"]), names, g["states"], g["ptn_freq"])
    return g, tree, model, aln


def check_nonrev(g, lib_path, prefix):
    """rooted tree (root = its last leaf), non-reversible model: lnL at the root branch, a leaf branch and internal
    branches, df/ddf where the branch has a length"""
    from tests.common import nonrev_engine, nonrev_eval
    from tests.test_nonrev import leaf_set  # noqa: F401
    ntaxa = int(g["ntaxa"])
    adj = [[int(v) for v in row if v >= 0] for row in g["adj_pad"]]
    length = {(int(a), int(b)): float(t) for a, b, t in g["blen"]}
    tree = W.Tree(ntaxa=ntaxa + 1, adj=adj, length=length, names=["T%d" % i for i in range(ntaxa + 1)])
    aln = W.Alignment(int(g["nstates"]), tree.names[:ntaxa], g["states"], g["ptn_freq"])
    nm = W.NonrevModel(nstates=int(g["nstates"]), Q=g["Q"], ceval=g["ceval"], cevec=g["cevec"], cinv_evec=g["cinv_evec"],
                       freqs=g["freqs"], mixw=g["mixw"])
    eng = nonrev_engine(aln, nm, g["rates"], g["props"], nslots=4 * ntaxa, rooted=True, lib_path=lib_path, prefix=prefix)
    for a, b, lnl, df, ddf in g["cases"]:
        a, b = int(a), int(b)
        got = nonrev_eval(eng, tree, a, b, ntaxa)
        assert rel(got["lnl"], lnl) <= TOL, (a, b)
        if a != ntaxa:
            assert rel(got["df"], df) <= TOL and rel(got["ddf"], ddf) <= TOL, (a, b)
    assert rel(got["pattern_lh"], g["pattern_lh"]) <= TOL


def check(make_engine, g, tree, model, aln):
    kind = str(g["kind"])
    if kind == "site":
        from tests.common import site_engine
        lib_path, prefix = make_engine("lib")
        eng = site_engine(aln, (g["s_eval"], g["s_evec"], g["s_inv"]), g["rates"], g["props"], nslots=tree.ntaxa,
                          safe_numeric=bool(g["safe"]), lib_path=lib_path, prefix=prefix)
        c2 = tree.adj[2][0]
        tr2 = W.traversal(tree, c2, 2)
        eng.update_partials(tr2)
        d2, n2, L2 = (LEAF, 2), (INTERNAL, tr2[-1][2]), tree.get_length(2, c2)
        assert rel(eng.branch_lnl(d2, n2, L2), g["lnl"]) <= TOL
        assert rel(eng.fetch_pattern_lh(), g["pattern_lh"]) <= TOL
        eng.build_theta(d2, n2)
        assert rel(eng.derv(L2), g["derv"]) <= TOL and rel(eng.lnl_from_buffer(L2), g["lnl_buf"]) <= TOL
        assert rel(eng.derv(2.5 * L2), g["derv2"]) <= TOL and rel(eng.lnl_from_buffer(2.5 * L2), g["lnl_buf2"]) <= TOL
        return
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    dad, node = (LEAF, 0), (INTERNAL, tr[-1][2])
    L = tree.get_length(0, c)
    if kind == "mixlen":
        eng = make_engine(safe_numeric=bool(g["safe"]))
        mult = g["mult"]
        eng.update_partials_mixlen(tr, mix_child_len(tr, mult))
        lens = L * mult
        assert rel(eng.branch_lnl_mixlen(dad, node, lens), g["lnl"]) <= TOL
        assert rel(eng.fetch_pattern_lh(), g["pattern_lh"]) <= TOL
        cat, want = eng.fetch_pattern_lh_cat(), g["pattern_lh_cat"]
        d = np.maximum(np.abs(want).max(axis=1, keepdims=True), 1e-300)
        assert float(np.max(np.abs(cat - want) / d)) <= TOL
        eng.build_theta(dad, node)
        df, ddf, lnl = eng.derv_mixlen(lens)
        assert rel(df, g["df"]) <= TOL and rel(ddf, g["ddf"]) <= TOL and rel(lnl, g["lnl_derv"]) <= TOL
        assert rel(eng.lnl_from_buffer_mixlen(lens * 1.25), g["lnl_buf"]) <= TOL
        for k in range(len(mult)):
            assert rel(eng.derv_mixlen_class(k, lens[k] * 0.8), g["cls"][k]) <= TOL
    elif kind == "holder":
        eng = make_engine(safe_numeric=bool(g["safe"]), unobserved=holder_unobserved(aln), asc_holder=True)
        eng.update_partials(tr)
        assert rel(eng.branch_lnl(dad, node, L), g["lnl"]) <= TOL
        assert rel(eng.fetch_pattern_lh(), g["pattern_lh"]) <= TOL
        eng.build_theta(dad, node)
        assert rel(eng.derv(L), g["derv"]) <= TOL
        assert rel([eng.lnl_from_buffer(t) for t in (0.5 * L, 2.0 * L)], g["buf"]) <= TOL
        assert rel(eng.derv(1.7 * L), g["derv2"]) <= TOL
    else:
        eng = make_engine()
        eng.update_partials(tr)
        for (keep, med), want in zip(g["settings"], g["lnl"]):
            eng.set_robust(float(keep), bool(med))
            assert rel(eng.branch_lnl(dad, node, L), want) <= TOL, (keep, med)


def test_variant_golden_files_present():
    assert len(FILES) >= 9


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[4:-4] for f in FILES])
def test_oracle_matches_reference_variant_golden(oracle_lib, path):
    import oracle
    g, tree, model, aln = load(path)
    oracle.set_ref_quirks(True)
    if str(g["kind"]) == "nonrev":
        return check_nonrev(g, oracle.ORACLE_LIB, "orc_")

    def make(*a, **kw):
        if a:
            return oracle.ORACLE_LIB, "orc_"
        return oracle.oracle_engine(aln, model, g["rates"], g["props"], nslots=tree.ntaxa, **kw)
    check(make, g, tree, model, aln)


@pytest.mark.gpu
@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[4:-4] for f in FILES])
def test_cuda_matches_reference_variant_golden(path):
    from iqtree2_b200 import capi
    g, tree, model, aln = load(path)
    if str(g["kind"]) == "nonrev":
        return check_nonrev(g, capi.LIB_PATH, "b200lk_")

    def make(*a, **kw):
        if a:
            return capi.LIB_PATH, "b200lk_"
        return engine_for(aln, model, g["rates"], g["props"], nslots=tree.ntaxa, **kw)
    check(make, g, tree, model, aln)
