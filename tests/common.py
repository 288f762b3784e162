"""Shared helpers for the parity tests: seeded problems and the call sequence
that is run identically against the CUDA engine and the CPU oracle."""
import numpy as np

from iqtree2_b200 import workload as W
from iqtree2_b200.capi import INTERNAL, LEAF


def make_problem(ntaxa, nsite, nstates=4, ncat=4, nmix=1, seed=1, alpha=0.5, compress=True,
                 gaps=0.0, mean_len=0.1):
    tree = W.random_tree(ntaxa, seed, mean_len=mean_len)
    if nstates == 4 and nmix == 1:
        model = W.gtr_model()
    else:
        model = W.random_reversible_model(nstates, seed + 100, nmix)
    rates, props = W.discrete_gamma(alpha, ncat)
    aln = W.simulate_alignment(tree, model, rates, props, nsite, seed + 7, compress=compress)
    if gaps > 0:
        rng = np.random.default_rng(seed + 13)
        mask = rng.random(aln.states.shape) < gaps
        amb = rng.integers(nstates, aln.state_unknown + 1, size=aln.states.shape).astype(np.uint8)
        aln.states[mask] = amb[mask]
    return tree, model, rates, props, aln


def full_eval(eng, tree, leaf=0, slot_base=0):
    """All partials stale -> traversal toward `leaf`, lnL at its branch, theta, df/ddf,
    lnL from buffer.  Returns a dict of everything comparable."""
    c = tree.adj[leaf][0]
    tr = W.traversal(tree, c, leaf)
    if slot_base:
        tr = [(a, b, s + slot_base, [(k, (i + slot_base if k == INTERNAL else i), l) for k, i, l in ch])
              for a, b, s, ch in tr]
    eng.update_partials(tr)
    L = tree.get_length(leaf, c)
    top = tr[-1][2]
    out = {"traversal": tr, "top": top}
    out["lnl"] = eng.branch_lnl((LEAF, leaf), (INTERNAL, top), L)
    out["pattern_lh"] = eng.fetch_pattern_lh()
    out["pattern_lh_cat"] = eng.fetch_pattern_lh_cat()
    eng.build_theta((LEAF, leaf), (INTERNAL, top))
    out["df"], out["ddf"] = eng.derv(L)
    out["lnl_buf"] = eng.lnl_from_buffer(L)
    out["theta"], out["bscale"] = eng.fetch_theta()
    out["scale"] = {s: eng.fetch_scale_num(s) for (_, _, s, _) in tr}
    out["partial_top"] = eng.fetch_partial(top)
    return out


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.maximum(np.abs(a), np.abs(b))))) if a.size else 0.0


def assert_parity(got, want, tol=1e-9):
    """north_star tolerance: scale counts bit-exact; lnL, df, ddf, per-pattern values 1e-9 relative."""
    for s in want["scale"]:
        assert np.array_equal(got["scale"][s], want["scale"][s]), "scale_num differs in slot %d" % s
    for k in ("lnl", "lnl_buf", "df", "ddf"):
        assert rel(got[k], want[k]) <= tol, (k, got[k], want[k])
    for k in ("pattern_lh", "bscale"):
        assert rel(got[k], want[k]) <= tol, k
    # values that underflow toward 0 are compared relative to the row's largest magnitude
    for k in ("theta", "partial_top", "pattern_lh_cat"):
        g, w = got[k].reshape(got[k].shape[0], -1), want[k].reshape(want[k].shape[0], -1)
        denom = np.maximum(np.abs(w).max(axis=1, keepdims=True), 1e-300)
        assert float(np.max(np.abs(g - w) / denom)) <= tol, k


def site_models_for(aln, nstates, seed, exch=None):
    """PMSF-shaped site-specific models: one exchangeability matrix, a frequency vector per PATTERN.
    -> (eval [nptn][n], evec [nptn][n][n], inv_evec [nptn][n][n])"""
    rng = np.random.default_rng(seed)
    n = nstates
    if exch is None:
        R = np.zeros((n, n))
        R[np.triu_indices(n, 1)] = rng.gamma(2.0, 1.0, n * (n - 1) // 2)
        exch = R + R.T
    ev, U, Ui = [], [], []
    for _ in range(aln.nptn):
        fr = rng.dirichlet(np.full(n, 3.0))
        Q, pi = W.reversible_model(exch, fr)
        w, evec, inv = W._decompose(Q, pi)
        ev.append(w); U.append(evec); Ui.append(inv)
    return np.ascontiguousarray(ev), np.ascontiguousarray(U), np.ascontiguousarray(Ui)


def site_engine(aln, models, rates, props, *, nslots, safe_numeric=False, lib_path=None, prefix="b200lk_", p_invar=0.0):
    """A context in site-specific mode (b200lk_dims.site_specific) with the per-pattern eigen-systems loaded."""
    from iqtree2_b200 import capi
    eng = capi.Engine(ntaxa=aln.ntaxa, nstates=aln.nstates, ncat=len(rates), nmix=1, nptn=aln.nptn, nslots=nslots,
                      n_state_codes=aln.state_unknown + 1, safe_numeric=safe_numeric, site_specific=True,
                      lib_path=lib_path or capi.LIB_PATH, prefix=prefix)
    eng.upload_tips(aln.states, aln.ptn_freq)
    eng.set_ptn_invar(np.zeros(aln.nptn))
    eng.set_site_models(models[0], models[1], models[2], rates, np.asarray(props) * (1.0 - p_invar),
                        W.state_indicators(aln.nstates))
    return eng


# ---- non-reversible kernel family (tree/phylokernelnonrev.h) ---------------------------------------------------------
class _RefModelShim:
    """What oracle.RefRun reads from a model object, for a NonrevModel (the eigen-system fields are placeholders: the
    non-reversible kernels never read them)."""

    def __init__(self, nm):
        n, k = nm.nstates, nm.nmix
        self.nmix = k
        self.eval = np.zeros((k, n))
        self.evec = np.tile(np.eye(n), (k, 1, 1))
        self.inv_evec = np.tile(np.eye(n), (k, 1, 1))
        self.mixw, self.freqs = nm.mixw, nm.freqs


def nonrev_engine(aln, nm, rates, props, *, nslots, rooted, safe_numeric=False, lib_path=None, prefix="b200lk_",
                  p_invar=0.0, ptn_invar=None):
    """A context of the non-reversible kernel family with the model loaded."""
    from iqtree2_b200 import capi
    eng = capi.Engine(ntaxa=aln.ntaxa, nstates=aln.nstates, ncat=len(rates), nmix=nm.nmix, nptn=aln.nptn, nslots=nslots,
                      n_state_codes=aln.state_unknown + 1, safe_numeric=safe_numeric, nonrev=True,
                      lib_path=lib_path or capi.LIB_PATH, prefix=prefix)
    eng.upload_tips(aln.states, aln.ptn_freq)
    eng.set_ptn_invar(np.zeros(aln.nptn) if ptn_invar is None else ptn_invar)
    eng.set_model_nonrev(nm.Q, nm.ceval, nm.cevec, nm.cinv_evec, nm.freqs, W.state_indicators(aln.nstates), rates,
                         np.asarray(props) * (1.0 - p_invar), nm.mixw, rooted)
    return eng


def nonrev_ends(tree, a, b, root_leaf):
    """The two ends of branch (a, b) in the order the non-reversible kernels use after their swap
    (tree/phylokernelnonrev.h:1060-1067): -> (dad, node); dad is the leaf end if there is one (the root leaf only when
    it is ``a``, the caller's dad), otherwise the end farther from the root.  ``root_leaf`` None: unrooted tree
    (params->kernel_nonrev), only the leaf rule applies."""
    dad, node = a, b
    if tree.is_leaf(node):
        dad, node = node, dad
    elif root_leaf is not None and not tree.is_leaf(dad):
        parent = tree.parents_from(root_leaf)
        if parent[node] == dad:      # dad_branch points AWAY from the root: swap
            dad, node = node, dad
    return dad, node


def nonrev_eval(eng, tree, a, b, root_leaf, want_derv=True):
    """All partials stale -> both traversals toward branch (a, b), lnL there, df/ddf, lnL again through the recorded
    ends.  Works on either backend (CUDA engine / oracle)."""
    from iqtree2_b200.capi import INTERNAL, LEAF, ROOT
    dad, node = nonrev_ends(tree, a, b, root_leaf)
    slot_of, entries = {}, []
    trav = (lambda v, u: W.traversal_rooted(tree, v, u, root_leaf, slot_of)) if root_leaf is not None else \
        (lambda v, u: W.traversal(tree, v, u, slot_of))
    # hand out slots up front so that both sides share one numbering
    for v, u in ((node, dad), (dad, node)):
        if not tree.is_leaf(v):
            for x, y, _, _ in W.traversal(tree, v, u):
                slot_of.setdefault((x, y), len(slot_of))
    for v, u in ((node, dad), (dad, node)):
        if not tree.is_leaf(v):
            entries += trav(v, u)
    eng.update_partials(entries)
    L = tree.get_length(a, b)
    node_end = (INTERNAL, slot_of[(node, dad)])
    if tree.is_leaf(dad):
        dad_end = (ROOT, 0) if dad == root_leaf else (LEAF, dad)
    else:
        dad_end = (INTERNAL, slot_of[(dad, node)])
    out = {"entries": entries, "ends": (dad_end, node_end)}
    out["lnl"] = eng.branch_lnl(dad_end, node_end, L)
    out["pattern_lh"] = eng.fetch_pattern_lh()
    out["pattern_lh_cat"] = eng.fetch_pattern_lh_cat()
    out["scale"] = {s: eng.fetch_scale_num(s) for (_, _, s, _) in entries}
    out["partials"] = {s: eng.fetch_partial(s) for (_, _, s, _) in entries}
    if want_derv and dad != root_leaf:
        eng.build_theta(dad_end, node_end)
        out["df"], out["ddf"] = eng.derv(L)
        out["lnl_buf"] = eng.lnl_from_buffer(L * 1.3)
    return out
