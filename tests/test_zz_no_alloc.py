"""b200lk_lnl_from_buffer_noalloc (include/b200lk.h): the from-buffer lnL computed out of the library's file-scope
`static __device__` scratch array instead of per-context cudaMalloc'ed buffers.  Checked through the C ABI against the
existing entry point b200lk_lnl_from_buffer, on calls in a row (the array is shared and keeps the last call's contents).

This file is named so that pytest collects it LAST."""
import numpy as np
import pytest

from iqtree2_b200 import workload as W
from iqtree2_b200.capi import INTERNAL, LEAF, B200LKError, Engine, engine_for

from .common import make_problem, rel

pytestmark = pytest.mark.gpu
TOL = 1e-9  # north_star tolerance for lnL and per-pattern lnL (relative)


def _prepared(nstates, nmix, ntaxa, nsite, seed, **kw):
    tree, model, rates, props, aln = make_problem(ntaxa, nsite, nstates=nstates, nmix=nmix, seed=seed, gaps=0.05)
    eng = engine_for(aln, model, rates, props, nslots=ntaxa, **kw)
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    eng.update_partials(tr)
    eng.build_theta((LEAF, 0), (INTERNAL, tr[-1][2]))
    return eng, tree.get_length(0, c)


@pytest.mark.parametrize("nstates,nmix,ntaxa,nsite,kw", [
    (4, 1, 24, 6000, {}),                       # T32 layout (4-state path)
    (4, 1, 24, 3000, {"safe_numeric": True}),   # SAFE_LH
    (20, 1, 12, 700, {}),                       # pattern-major layout, block = 80
    (20, 3, 10, 300, {}),                       # mixture: block = 240
    (4, 1, 16, 2000, {"p_invar": 0.2}),         # +I
])
def test_noalloc_matches_lnl_from_buffer_on_calls_in_a_row(nstates, nmix, ntaxa, nsite, kw):
    eng, L = _prepared(nstates, nmix, ntaxa, nsite, 11, **kw)
    # two different lengths, then the first again: the shared scratch still holds the previous call's tables and sums
    for length in (L, 2.5 * L + 0.01, L):
        want = eng.lnl_from_buffer(length)
        want_plh = eng.fetch_pattern_lh()
        got = eng.lnl_from_buffer_noalloc(length)
        got_plh = eng.fetch_pattern_lh()
        assert rel(got, want) <= TOL, (length, got, want)
        assert rel(got_plh, want_plh) <= TOL
        again = eng.lnl_from_buffer_noalloc(length)
        assert again == got  # deterministic reduction: bit-identical on a repeat
    eng.close()


def test_noalloc_interleaves_contexts_that_share_the_scratch():
    a, La = _prepared(4, 1, 20, 4000, 21)
    b, Lb = _prepared(20, 1, 10, 500, 22)
    wa, wb = a.lnl_from_buffer(La), b.lnl_from_buffer(Lb)
    for _ in range(2):  # serialised calls of two contexts alternate on the one array
        assert rel(a.lnl_from_buffer_noalloc(La), wa) <= TOL
        assert rel(b.lnl_from_buffer_noalloc(Lb), wb) <= TOL
    a.close()
    b.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch_and_wrong_state():
    # block = 64 states x 4 categories x 20 classes = 5120 doubles > B200LK_NOALLOC_MAX_BLOCK (4800)
    eng = Engine(ntaxa=4, nstates=64, ncat=4, nmix=20, nptn=8, nslots=4, n_state_codes=65)
    eng.upload_tips(np.zeros((4, 8), dtype=np.uint8), np.ones(8))
    with pytest.raises(B200LKError, match="-3"):  # no model, no theta yet
        eng.lnl_from_buffer_noalloc(0.1)
    eng.close()
    tree, model, rates, props, aln = make_problem(4, 40, nstates=64, nmix=20, seed=5)
    eng = engine_for(aln, model, rates, props, nslots=4)
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    eng.update_partials(tr)
    eng.build_theta((LEAF, 0), (INTERNAL, tr[-1][2]))
    assert np.isfinite(eng.lnl_from_buffer(0.1))  # the allocating entry point serves this size
    with pytest.raises(B200LKError, match="exceeds the static scratch"):
        eng.lnl_from_buffer_noalloc(0.1)
    eng.close()
