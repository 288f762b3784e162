"""Reversible eigen-decomposition: the oracle's restatement against the reference's own
EigenDecomposition::eigensystem_sym (oracle/_ref) and against numpy; the device solver against the oracle."""
import ctypes as C

import numpy as np
import pytest

import oracle
from iqtree2_b200 import workload as W


def random_rates(n, seed, zero_states=()):
    rng = np.random.default_rng(seed)
    exch = rng.gamma(2.0, 1.0, n * (n - 1) // 2)
    freqs = rng.dirichlet(np.full(n, 5.0))
    for z in zero_states:
        freqs[z] = 0.0
    return exch, freqs / freqs.sum()


def p_matrix(eval_, evec, inv, t):
    return (evec * np.exp(eval_ * t)) @ inv


def orc_eigen(exch, freqs, total=1.0):
    n = len(freqs)
    lib = C.CDLL(oracle.ORACLE_LIB)
    ev, U, Ui = np.zeros(n), np.zeros((n, n)), np.zeros((n, n))
    dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    lib.orc_eigen_rev.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_double] + [C.POINTER(C.c_double)] * 3
    assert lib.orc_eigen_rev(n, dp(np.ascontiguousarray(exch)), dp(np.ascontiguousarray(freqs)), total, dp(ev), dp(U), dp(Ui)) == 0
    return ev, U, Ui


def ref_eigen(exch, freqs, total=1.0):
    n = len(freqs)
    lib = C.CDLL(oracle.REF_LIB)
    ev, U, Ui = np.zeros(n), np.zeros((n, n)), np.zeros((n, n))
    dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    lib.ref_eigensystem_sym.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_double] + [C.POINTER(C.c_double)] * 3
    lib.ref_eigensystem_sym.restype = None
    lib.ref_eigensystem_sym(n, dp(np.ascontiguousarray(exch)), dp(np.ascontiguousarray(freqs)), total, dp(ev), dp(U), dp(Ui))
    return ev, U, Ui


def q_matrix(exch, freqs, total=1.0):
    n = len(freqs)
    R = np.zeros((n, n))
    R[np.triu_indices(n, 1)] = exch
    R = R + R.T
    Q = R * freqs[None, :]
    np.fill_diagonal(Q, 0.0)
    np.fill_diagonal(Q, -Q.sum(axis=1))
    return Q * (total / -(freqs * np.diag(Q)).sum())


@pytest.mark.parametrize("n,seed", [(4, 1), (20, 2), (61, 3), (5, 4), (2, 5)])
def test_oracle_eigen_reconstructs_rate_matrix(oracle_lib, n, seed):
    exch, freqs = random_rates(n, seed)
    ev, U, Ui = orc_eigen(exch, freqs)
    Q = q_matrix(exch, freqs)
    assert np.allclose((U * ev) @ Ui, Q, atol=1e-12)
    assert np.allclose(U @ Ui, np.eye(n), atol=1e-12)
    assert abs(max(ev)) < 1e-12 and (np.sort(ev)[:-1] < 0).all()
    for t in (0.05, 1.3):
        P = p_matrix(ev, U, Ui, t)
        assert np.allclose(P.sum(axis=1), 1.0, atol=1e-12) and (P > -1e-14).all()


@pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("n,seed,zero", [(4, 11, ()), (20, 12, ()), (61, 13, ()), (20, 14, (3, 17)), (7, 15, (0,))])
def test_oracle_eigen_matches_reference_solver(oracle_lib, n, seed, zero):
    """Same algorithm (tred2 + tqli on the symmetrised matrix): eigenvalues in the same order, eigenvectors up to
    rounding; and the quantities the kernels consume -- P(t) -- to 1e-13."""
    exch, freqs = random_rates(n, seed, zero)
    for total in (1.0, 0.37):
        eo, Uo, Io = orc_eigen(exch, freqs, total)
        er, Ur, Ir = ref_eigen(exch, freqs, total)
        assert np.allclose(eo, er, rtol=1e-11, atol=1e-13)
        for t in (0.01, 0.4, 5.0):
            assert np.allclose(p_matrix(eo, Uo, Io, t), p_matrix(er, Ur, Ir, t), atol=1e-13)
        # column signs aside, the eigenvectors themselves agree
        for j in range(n):
            s = 1.0 if np.dot(Uo[:, j], Ur[:, j]) >= 0 else -1.0
            assert np.allclose(Uo[:, j], s * Ur[:, j], atol=1e-9)


def test_workload_models_agree_with_oracle_eigen(oracle_lib):
    """The eigen-systems the parity tests feed the kernels (numpy eigh in workload.py) span the same P(t)."""
    exch = np.array([1.0, 2.0, 1.5, 1.2, 2.5, 1.0])
    freqs = np.array([0.3, 0.2, 0.2, 0.3])
    m = W.gtr_model(tuple(exch), tuple(freqs))
    ev, U, Ui = orc_eigen(exch, freqs)
    for t in (0.1, 2.0):
        assert np.allclose(p_matrix(m.eval[0], m.evec[0], m.inv_evec[0], t), p_matrix(ev, U, Ui, t), atol=1e-13)


# ---- the device solver (eigen.cu) against the oracle ---------------------------------------------------------------
from iqtree2_b200.capi import INTERNAL, LEAF, Batch, engine_for  # noqa: E402
from tests.common import assert_parity, full_eval, make_problem, rel  # noqa: E402


def _engines_from_rates(aln, exch, freqs, rates, props, mixw, nslots, total=None, safe=False):
    """(CUDA engine, oracle engine), both given only the model's free parameters."""
    nmix = len(mixw)
    dummy = W.random_reversible_model(aln.nstates, 1, nmix)
    tipvec = W.state_indicators(aln.nstates)
    out = []
    for mk in (engine_for, oracle.oracle_engine):
        e = mk(aln, dummy, rates, props, nslots=nslots, safe_numeric=safe)
        e.set_model_rates(exch, freqs, rates, props, mixw, tipvec, total)
        out.append(e)
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("n,nmix,zero", [(4, 1, ()), (20, 1, ()), (20, 3, ()), (61, 1, ()), (5, 2, ()), (2, 1, ()), (20, 1, (2, 11)),
                                         (33, 1, ())])
def test_device_eigen_matches_oracle(oracle_lib, n, nmix, zero):
    rng = np.random.default_rng(50 + n)
    tree, _, rates, props, aln = make_problem(6, 40, nstates=n, ncat=2, seed=3)
    ex, fr = zip(*[random_rates(n, 70 + n + k, zero) for k in range(nmix)])
    ex, fr = np.array(ex), np.array(fr)
    total = rng.uniform(0.3, 2.0, nmix)
    mixw = rng.dirichlet(np.ones(nmix))
    g, o = _engines_from_rates(aln, ex, fr, rates, props, mixw, 6, total)
    eg, Ug, Ig = g.fetch_model()
    for k in range(nmix):
        eo, Uo, Io = orc_eigen(ex[k], fr[k], total[k])
        assert np.allclose(np.sort(eg[k]), np.sort(eo), rtol=1e-11, atol=1e-13)
        assert np.allclose(Ug[k] @ Ig[k], np.eye(n), atol=1e-12)
        for t in (0.02, 0.7, 4.0):
            assert np.allclose(p_matrix(eg[k], Ug[k], Ig[k], t), p_matrix(eo, Uo, Io, t), atol=2e-13)


@pytest.mark.gpu
@pytest.mark.parametrize("n,nmix,safe", [(4, 1, False), (4, 2, True), (20, 1, False), (20, 2, False), (12, 1, False), (6, 1, True), (61, 1, False)])
def test_likelihood_from_model_parameters(oracle_lib, n, nmix, safe):
    """lnL, df, ddf, scale counts with the eigen-system solved on the device = the oracle's with the reference's solver."""
    ntaxa = 9
    tree, _, rates, props, aln = make_problem(ntaxa, 160, nstates=n, ncat=3, nmix=nmix, seed=23, gaps=0.05)
    ex, fr = zip(*[random_rates(n, 200 + n + k) for k in range(nmix)])
    mixw = np.full(nmix, 1.0 / nmix)
    oracle.set_ref_quirks(False)
    g, o = _engines_from_rates(aln, np.array(ex), np.array(fr), rates, props, mixw, ntaxa, None, safe)
    rg, ro = full_eval(g, tree, 1), full_eval(o, tree, 1)
    for k in ("lnl", "lnl_buf", "df", "ddf"):
        assert rel(rg[k], ro[k]) <= 1e-9, k
    assert rel(rg["pattern_lh"], ro["pattern_lh"]) <= 1e-9
    for s in ro["scale"]:
        assert np.array_equal(rg["scale"][s], ro["scale"][s])


DNA_FAMILY = {  # exchangeabilities AC AG AT CG CT GT of the nested DNA models ModelFinder walks through
    "JC": ((1, 1, 1, 1, 1, 1), False), "K80": ((1, 3.1, 1, 1, 3.1, 1), False), "HKY": ((1, 3.1, 1, 1, 3.1, 1), True),
    "TN": ((1, 2.4, 1, 1, 4.2, 1), True), "SYM": ((1.3, 3.1, 0.7, 1.1, 4.0, 1), False), "GTR": ((1.3, 3.1, 0.7, 1.1, 4.0, 1), True),
}


@pytest.mark.gpu
def test_batched_model_probes_on_one_alignment(oracle_lib):
    """ModelFinder's inner loop: several candidate models on the SAME alignment and tree, eigen-systems solved on the
    device, all candidates evaluated by one batched launch sequence (b200lk_batch_*)."""
    ntaxa = 16
    tree, _, rates, props, aln = make_problem(ntaxa, 900, seed=77, gaps=0.03)
    emp = np.array([(aln.states == s).sum() for s in range(4)], dtype=float)
    emp /= emp.sum()
    engines, oracles = [], []
    for name, (ex, use_freq) in DNA_FAMILY.items():
        fr = emp if use_freq else np.full(4, 0.25)
        g, o = _engines_from_rates(aln, np.array([ex], dtype=float), np.array([fr]), rates, props, np.ones(1), ntaxa)
        engines.append(g)
        oracles.append(o)
    leaf = 0
    c = tree.adj[leaf][0]
    tr = W.traversal(tree, c, leaf)
    top = tr[-1][2]
    nch = sum(len(e[3]) for e in tr)
    child_len = np.tile(np.array([ch[2] for e in tr for ch in e[3]]), (len(engines), 1))
    blen = np.full(len(engines), tree.get_length(leaf, c))
    assert child_len.shape == (len(engines), nch)
    lnl, df, ddf = Batch(engines).evaluate(tr, child_len, (LEAF, leaf), (INTERNAL, top), blen)
    want = [full_eval(o, tree, leaf) for o in oracles]
    for k in range(len(engines)):
        assert rel(lnl[k], want[k]["lnl"]) <= 1e-9 and rel(df[k], want[k]["df"]) <= 1e-9 and rel(ddf[k], want[k]["ddf"]) <= 1e-9
    # nested models: more free parameters never fit worse at the same branch lengths... not guaranteed without
    # optimisation, but the candidates must at least differ
    assert len({round(float(x), 6) for x in lnl}) == len(engines)
