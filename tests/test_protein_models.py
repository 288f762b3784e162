"""LG and LG+C60 as data (BASELINE configs 3 and 4): the tables extracted by tools/extract_model_data.py,
the models built from them in Python (workload.lg_model / lg_c60_model) and in the C++ host layer (ModelSubst::LG /
LGC60), pinned against the reference's own eigen-solver and, through the golden vectors the reference's kernel wrote
(tests/golden/ref_aa_lg_g4.npz, ref_aa_lg_c60_g4.npz), against the reference's likelihood."""
import json
import os
import sys

import numpy as np
import pytest

import oracle
from iqtree2_b200 import build_host, host as H, workload as W
from iqtree2_b200.capi import INTERNAL, LEAF

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HERE = os.path.dirname(os.path.abspath(__file__))


def p_of(ev, U, Ui, t):
    return np.einsum("mxi,mi,miy->mxy", U, np.exp(ev * t), Ui)


def test_tables_have_the_published_shape():
    d = W.protein_data()
    lg, c60 = d["LG"], d["C60"]
    assert d["states"] == "ARNDCQEGHILKMFPSTWYV"
    assert len(lg["lower_triangle"]) == 190 and len(lg["freqs"]) == 20
    assert abs(sum(lg["freqs"]) - 1.0) < 1e-6
    assert lg["lower_triangle"][0] == 0.425093 and lg["lower_triangle"][-1] == 0.249313  # R<->A, V<->Y
    assert len(c60["profiles"]) == 60 and all(len(p) == 20 for p in c60["profiles"]) and len(c60["weights"]) == 60
    assert abs(sum(c60["weights"]) - 1.0) < 1e-8
    assert max(abs(sum(p) - 1.0) for p in c60["profiles"]) < 1e-6
    assert c60["weights"][45] == 0.0004150083  # C60pi46, the rarest class


@pytest.mark.skipif(not os.path.isdir("/root/reference/model"), reason="the reference checkout is not on this machine")
def test_tables_equal_the_reference_files():
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import extract_model_data as X
    fresh = X.parse("/root/reference")
    assert json.loads(json.dumps(fresh)) == W.protein_data()


def test_models_are_reversible_and_normalised():
    for m in (W.lg_model(), W.lg_c60_model()):
        for k in range(m.nmix):
            Q, pi = m.Q[k], m.freqs[k]
            assert np.allclose(Q.sum(axis=1), 0.0, atol=1e-12)
            assert np.allclose(pi[:, None] * Q, (pi[:, None] * Q).T, atol=1e-14)  # detailed balance
            assert abs(-(pi * np.diag(Q)).sum() - 1.0) < 1e-12                   # one expected substitution per class
            assert np.allclose(m.evec[k] @ m.inv_evec[k], np.eye(20), atol=1e-11)
        assert abs(m.mixw.sum() - 1.0) < 1e-12
    assert W.lg_c60_model().nmix == 60


def test_host_layer_models_equal_the_python_ones():
    build_host.build()
    for name, m in (("LG", W.lg_model()), ("LG+C60", W.lg_c60_model())):
        ev, U, Ui, fr, w = H.HostModel.named(name).eigen()
        assert np.abs(fr - m.freqs).max() < 1e-15 and np.abs(w - m.mixw).max() < 1e-15
        for t in (0.01, 0.37, 2.5):
            assert np.abs(p_of(ev, U, Ui, t) - p_of(m.eval, m.evec, m.inv_evec, t)).max() < 1e-12


@pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref (the reference's own solver) is not built here")
def test_eigen_systems_match_the_reference_solver():
    from tests.test_eigen import ref_eigen
    exch = W.lg_exchangeabilities()[np.triu_indices(20, 1)]
    for m in (W.lg_model(), W.lg_c60_model()):
        for k in range(0, m.nmix, 7):
            e, u, ui = ref_eigen(exch, m.freqs[k], 1.0)
            assert np.abs(np.sort(e) - np.sort(m.eval[k])).max() < 1e-11
            for t in (0.05, 1.3):
                assert np.abs(p_of(e[None], u[None], ui[None], t)[0] - p_of(m.eval[k:k + 1], m.evec[k:k + 1], m.inv_evec[k:k + 1], t)[0]).max() < 1e-12


@pytest.mark.parametrize("case,build", [("aa_lg_g4", W.lg_model), ("aa_lg_c60_g4", W.lg_c60_model)])
def test_own_model_construction_reproduces_the_reference_likelihood(oracle_lib, case, build):
    """The golden files carry the eigen-system of the REFERENCE's solver and what the reference's kernel computed with
    it.  The same alignment and tree under OUR model construction (numpy symmetric solve) must give the same lnL,
    df and ddf: this pins data + construction, not only the kernel."""
    g = np.load(os.path.join(HERE, "golden", "ref_%s.npz" % case), allow_pickle=False)
    names = [str(x) for x in g["names"]]
    tree = W.Tree.from_newick(str(g["newick"]), names)
    aln = W.Alignment(20, names, g["states"], g["ptn_freq"])
    model = build()
    assert np.abs(model.freqs - g["freqs"]).max() < 1e-15 and np.abs(model.mixw - g["mixw"]).max() < 1e-15
    eng = oracle.oracle_engine(aln, model, g["rates"], g["props"], nslots=tree.ntaxa)
    leaf = int(g["leaf"])
    c = tree.adj[leaf][0]
    tr = W.traversal(tree, c, leaf)
    lnl, df, ddf = eng.evaluate(tr, (LEAF, leaf), (INTERNAL, tr[-1][2]), tree.get_length(leaf, c))
    rel = lambda a, b: abs(a - b) / max(1.0, abs(a), abs(b))
    assert rel(lnl, float(g["lnl"])) <= 1e-9 and rel(df, float(g["df"])) <= 1e-9 and rel(ddf, float(g["ddf"])) <= 1e-9
    assert np.max(np.abs(eng.fetch_pattern_lh() - g["pattern_lh"])) <= 1e-9 * max(1.0, np.abs(g["pattern_lh"]).max())
