"""The adapter compiled against the REAL reference PhyloTree (adapter/Makefile: the reference's own tree/phylotree.cpp,
phylotreesse.cpp, memslot.cpp, utils/optimization.cpp ... + adapter/lk_cuda.patch + adapter/phylokernelcuda.cpp).
The reference's OWN callers -- computeLikelihood (tree/phylotree.cpp:1227-1268), computeTraversalInfo (:5907-6015),
reorientPartialLh (:5887-5901), optimizeOneBranch / minimizeNewton (:2585-2635), optimizeAllBranches (:2689-2765) and
getBestNNIForBran (:4058-4288, nni_partial_lh -> slots nslots-2 / nslots-1) -- run over libb200lk.so with `-lk CUDA` and
are compared with the same callers over the reference's AVX+FMA kernel (`-lk FMA`).  Bar: lnL 1e-9 relative, optimised
branch lengths 1e-6, the same NNI winner."""
import os

import numpy as np
import pytest

import oracle
from iqtree2_b200 import workload as W
from tests.common import make_problem, rel
from tests.test_oracle_vs_ref import leaf_sets

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not oracle.adapter_available(), reason="adapter/_build/libiqtree_cuda.so is not built")]
HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-9


def pair(tree, aln, model, rates, props, **kw):
    fma = oracle.RefRun(tree, aln, model, rates, props, kernel=oracle.LK_AVX_FMA, lib_path=oracle.ADAPTER_LIB, **kw)
    cuda = oracle.RefRun(tree, aln, model, rates, props, kernel=oracle.LK_CUDA, lib_path=oracle.ADAPTER_LIB, **kw)
    return cuda, fma


def example_phy():
    aln = W.alignment_from_phylip(os.path.join(HERE, "golden", "example.phy"))
    tree = W.Tree.from_newick(open(os.path.join(HERE, "golden", "example17.nwk")).read(), aln.names)
    rates, props = W.discrete_gamma(0.5)
    return tree, W.gtr_model(), rates, props, aln


def splits_of(tree):
    ls = leaf_sets(tree, W.traversal(tree, tree.adj[0][0], 0))
    return [sorted(v) for v in ls.values()]


def branch_lengths(run, tree):
    out = [run.branch_length(s, 0) for s in splits_of(tree)]
    out += [run.branch_length([t], t) for t in range(tree.ntaxa)]
    return np.array(out)


CASES = {
    "example_phy": lambda: (example_phy(), {}),
    "dna_200x10k": lambda: (make_problem(200, 10_000, seed=41, compress=False), {}),
    "dna_safe_invar": lambda: (make_problem(24, 800, seed=42, gaps=0.05), {"safe": True, "p_invar": 0.2}),
    "aa_mix3": lambda: (make_problem(9, 150, nstates=20, nmix=3, seed=43), {}),
    "dna_deep": lambda: (make_problem(150, 120, seed=44, mean_len=0.6), {}),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_reference_callers_over_cuda_match_fma(case):
    (tree, model, rates, props, aln), kw = CASES[case]()
    cuda, fma = pair(tree, aln, model, rates, props, **kw)
    assert cuda.vector_size == 1
    # computeLikelihoodBranch at a leaf, all partials stale; per-pattern values; df/ddf; from-buffer
    assert rel(cuda.lnl_at_leaf(0), fma.lnl_at_leaf(0)) <= TOL
    assert rel(cuda.pattern_lh(), fma.pattern_lh()) <= TOL
    assert rel(cuda.derv(), fma.derv()) <= TOL
    assert rel(cuda.from_buffer(), fma.from_buffer()) <= TOL
    # the reference's own root choice (computeLikelihood) and an internal branch
    assert rel(cuda.compute_likelihood(), fma.compute_likelihood()) <= TOL
    internal = [s for s in splits_of(tree) if 2 <= len(s) <= tree.ntaxa - 3]
    if internal:
        s = internal[len(internal) // 2]
        assert rel(cuda.lnl_at_branch(s, 0), fma.lnl_at_branch(s, 0)) <= TOL
    # optimizeAllBranches: the reference's loop, re-rooting (reorientPartialLh) and Newton-Raphson over the CUDA kernel
    if case != "dna_deep":
        lc, lf = cuda.optimize_all_branches(), fma.optimize_all_branches()
        assert rel(lc, lf) <= TOL
        assert np.max(np.abs(branch_lengths(cuda, tree) - branch_lengths(fma, tree))) <= 1e-6
    cuda.close()
    fma.close()


@pytest.mark.parametrize("nni5", [False, True])
@pytest.mark.parametrize("case", ["example_phy", "dna_200x10k", "aa_mix3"])
def test_nni_evaluation_over_cuda_matches_fma(case, nni5):
    """getBestNNIForBran: partial_lh of the two central directions re-pointed into nni_partial_lh (device slots
    nslots-2, nslots-1), optimizeOneBranch x1 (x5 with nni5), computeLikelihoodFromBuffer, everything restored."""
    (tree, model, rates, props, aln), kw = CASES[case]()
    cuda, fma = pair(tree, aln, model, rates, props, **kw)
    internal = [s for s in splits_of(tree) if 2 <= len(s) <= tree.ntaxa - 3]
    for s in internal[:: max(1, len(internal) // 4)][:4]:
        wc, llc, lnc, bc, ac = cuda.nni(s, 0, nni5)
        wf, llf, lnf, bf, af = fma.nni(s, 0, nni5)
        assert rel(bc, bf) <= TOL
        assert rel(llc, llf) <= 1e-8, (s, llc, llf)   # scores after Newton-Raphson on up to five branches
        assert np.max(np.abs(lnc - lnf)) <= 1e-6
        assert wc == wf
        assert rel(ac, bc) <= TOL and rel(ac, af) <= TOL   # the tree, its flags and its slots are restored
    cuda.close()
    fma.close()
