"""Golden vectors computed by the REFERENCE's own AVX+FMA kernels (tests/golden/make_golden.py, run where
/root/reference exists) checked against (a) the CPU oracle -- which pins the oracle on machines that do not
have the reference -- and (b) the CUDA path through the C ABI.  Bar: scale counts bit-exact; lnL, df, ddf,
per-pattern values within 1e-9 relative."""
import glob
import os

import numpy as np
import pytest

from iqtree2_b200 import workload as W
from iqtree2_b200.capi import INTERNAL, LEAF, engine_for

HERE = os.path.dirname(os.path.abspath(__file__))
FILES = sorted(glob.glob(os.path.join(HERE, "golden", "ref_*.npz")))
TOL = 1e-9


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.maximum(np.abs(a), np.abs(b)))))


def load(path):
    g = np.load(path, allow_pickle=False)
    names = [str(x) for x in g["names"]]
    tree = W.Tree.from_newick(str(g["newick"]), names)
    model = W.EigenModel(nstates=int(g["nstates"]), eval=g["eval"], evec=g["evec"], inv_evec=g["inv_evec"],
                         freqs=g["freqs"], mixw=g["mixw"], Q=np.zeros(0))
    aln = W.Alignment(int(g["nstates"]), names, g["states"], g["ptn_freq"])
    return g, tree, model, aln


def check(eng, g, tree):
    leaf = int(g["leaf"])
    c = tree.adj[leaf][0]
    tr = W.traversal(tree, c, leaf)
    eng.update_partials(tr)
    L = tree.get_length(leaf, c)
    top = tr[-1][2]
    assert rel(eng.branch_lnl((LEAF, leaf), (INTERNAL, top), L), g["lnl"]) <= TOL
    assert rel(eng.fetch_pattern_lh(), g["pattern_lh"]) <= TOL
    cat, want = eng.fetch_pattern_lh_cat(), g["pattern_lh_cat"]
    d = np.maximum(np.abs(want).max(axis=1, keepdims=True), 1e-300)
    assert float(np.max(np.abs(cat - want) / d)) <= TOL
    nz = 0
    for (_, _, slot, _) in tr:
        sc = g["scale_%d" % slot]
        assert np.array_equal(eng.fetch_scale_num(slot), sc), "scale_num differs from the reference, slot %d" % slot
        nz += int(sc.max() > 0)
        key = "partial_%d" % slot
        if key in g:
            want = g[key]
            got = eng.fetch_partial(slot)
            d = np.maximum(np.abs(want).reshape(want.shape[0], -1).max(axis=1), 1e-300)[:, None, None]
            assert float(np.max(np.abs(got - want) / d)) <= TOL
    eng.build_theta((LEAF, leaf), (INTERNAL, top))
    assert rel(eng.derv(L), (g["df"], g["ddf"])) <= TOL
    assert rel(eng.lnl_from_buffer(L), g["lnl_buf"]) <= TOL
    for t, dd, ll in zip(g["other_t"], g["other_derv"], g["other_lnl"]):
        assert rel(eng.derv(float(t)), dd) <= TOL
        assert rel(eng.lnl_from_buffer(float(t)), ll) <= TOL
    return nz


def test_golden_files_present():
    assert len(FILES) >= 7


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[4:-4] for f in FILES])
def test_oracle_matches_reference_golden(oracle_lib, path):
    import oracle
    g, tree, model, aln = load(path)
    oracle.set_ref_quirks(True)
    eng = oracle.oracle_engine(aln, model, g["rates"], g["props"], nslots=tree.ntaxa, safe_numeric=bool(g["safe"]),
                               p_invar=float(g["p_invar"]))
    nz = check(eng, g, tree)
    if "deep" in path:
        assert nz > 0


@pytest.mark.gpu
@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[4:-4] for f in FILES])
def test_cuda_matches_reference_golden(path):
    g, tree, model, aln = load(path)
    eng = engine_for(aln, model, g["rates"], g["props"], nslots=tree.ntaxa, safe_numeric=bool(g["safe"]),
                     p_invar=float(g["p_invar"]))
    nz = check(eng, g, tree)
    if "deep" in path:
        assert nz > 0
