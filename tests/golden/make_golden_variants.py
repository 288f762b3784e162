#!/usr/bin/env python
"""Generate tests/golden/var_*.npz: what the REFERENCE's own kernels (oracle/_ref) compute for the variant paths --
heterotachy (per-class branch lengths, the isMixlen() derivative, computeLikelihoodDervMixlen), Holder's +ASC and
robust-phylogeny scoring.  Run in the authoring container only (needs /root/reference):

    python tests/golden/make_golden_variants.py

Each file holds the complete inputs and the reference's outputs; tests/test_golden_variants.py checks the oracle (CPU)
and the CUDA path (GPU) against them where /root/reference does not exist."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle  # noqa: E402
from tests.common import make_problem  # noqa: E402
from tests.common import _RefModelShim, site_models_for  # noqa: E402
from tests.test_mixlen_holder import holder_problem, holder_unobserved, mixlen_problem  # noqa: E402
from tests.test_nonrev import internal_branches, leaf_set, rooted_problem  # noqa: E402
from iqtree2_b200 import workload as W  # noqa: E402


def base(tree, model, rates, props, aln, **extra):
    out = dict(nstates=aln.nstates, states=aln.states, ptn_freq=aln.ptn_freq, newick=tree.to_newick(0, 17),
               names=np.array(tree.names), eval=model.eval, evec=model.evec, inv_evec=model.inv_evec,
               freqs=model.freqs, mixw=model.mixw, rates=rates, props=props)
    out.update(extra)
    return out


def save(name, out):
    path = os.path.join(HERE, "var_%s.npz" % name)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


def main():
    # heterotachy
    for name, (n, k, safe) in {"mixlen_dna4": (4, 4, False), "mixlen_aa3_safe": (20, 3, True)}.items():
        tree, model, rates, props, aln, mult = mixlen_problem(n, k, seed=301 + n)
        ref = oracle.RefRun(tree, aln, model, rates, props, kernel=oracle.LK_AVX_FMA, safe=safe, mixlen=mult)
        out = base(tree, model, rates, props, aln, kind="mixlen", mult=mult, safe=safe)
        out["lnl"] = ref.lnl_at_leaf(0)
        out["pattern_lh"] = ref.pattern_lh()
        out["pattern_lh_cat"] = ref.pattern_lh_cat()
        df, ddf, lnl = ref.derv_mixlen(k)
        out.update(df=df, ddf=ddf, lnl_derv=lnl)
        lens = tree.get_length(0, tree.adj[0][0]) * mult
        ref.set_current_mixlen(lens * 1.25)
        out["lnl_buf"] = ref.from_buffer()
        ref.set_current_mixlen(lens)
        out["cls"] = np.array([ref.derv_mixlen_class(c, lens[c] * 0.8, again=True) for c in range(k)])
        ref.close()
        save(name, out)
    # Holder's +ASC
    for name, (n, safe) in {"holder_dna": (4, False), "holder_aa_safe": (20, True)}.items():
        tree, model, rates, props, aln = holder_problem(n, seed=311 + n)
        ref = oracle.RefRun(tree, aln, model, rates, props, kernel=oracle.LK_AVX_FMA, safe=safe, asc_holder=True,
                            unobserved=holder_unobserved(aln, V=4))
        out = base(tree, model, rates, props, aln, kind="holder", safe=safe)
        L = tree.get_length(0, tree.adj[0][0])
        out["lnl"] = ref.lnl_at_leaf(0)
        out["pattern_lh"] = ref.pattern_lh()
        out["derv"] = np.array(ref.derv())
        buf = []
        for t in (0.5 * L, 2.0 * L):
            ref.set_current_length(t)
            buf.append(ref.from_buffer())
        out["buf"] = np.array(buf)
        ref.set_current_length(1.7 * L)
        out["derv2"] = np.array(ref.derv(again=True))
        ref.close()
        save(name, out)
    # robust phylogeny
    tree, model, rates, props, aln = make_problem(9, 301, nstates=4, seed=321)
    out = base(tree, model, rates, props, aln, kind="robust")
    settings = [(0.9, 0), (0.5, 0), (1.0, 1)]
    vals = []
    for keep, med in settings:
        ref = oracle.RefRun(tree, aln, model, rates, props, kernel=oracle.LK_AVX_FMA, robust=(keep, med))
        vals.append(ref.lnl_at_leaf(0))
        ref.close()
    out.update(settings=np.array(settings, dtype=np.float64), lnl=np.array(vals))
    save("robust_dna", out)
    # site-specific models (PMSF / ModelSet): SITE_MODEL kernel variants
    for name, (n, ncat, safe) in {"site_aa": (20, 4, False), "site_dna_safe": (4, 3, True)}.items():
        tree, _, rates, props, aln = make_problem(9, 61 if n == 20 else 150, nstates=n, ncat=ncat, seed=331, gaps=0.06)
        models = site_models_for(aln, n, 332)
        dummy = W.random_reversible_model(n, 1)
        ref = oracle.RefRun(tree, aln, dummy, rates, props, kernel=oracle.LK_AVX_FMA, safe=safe, site_models=models)
        out = base(tree, dummy, rates, props, aln, kind="site", safe=ref.safe_numeric, s_eval=models[0], s_evec=models[1],
                   s_inv=models[2])
        L = tree.get_length(2, tree.adj[2][0])
        out["lnl"] = ref.lnl_at_leaf(2)
        out["pattern_lh"] = ref.pattern_lh()
        out["derv"] = np.array(ref.derv())
        out["lnl_buf"] = ref.from_buffer()
        ref.set_current_length(2.5 * L)
        out["derv2"] = np.array(ref.derv(again=True))
        out["lnl_buf2"] = ref.from_buffer()
        ref.close()
        save(name, out)
    # non-reversible family on a rooted tree
    for name, (n, ncat, nmix) in {"nonrev_dna": (4, 4, 1), "nonrev_aa": (20, 2, 1)}.items():
        ntaxa = 9
        tree, nm, rates, props, aln = rooted_problem(ntaxa, 200, n, ncat, nmix, seed=341)
        ref = oracle.RefRun(tree, aln, _RefModelShim(nm), rates, props, safe=False,
                            nonrev=(1, nm.Q, nm.ceval, nm.cevec, nm.cinv_evec))
        out = dict(nstates=aln.nstates, states=aln.states, ptn_freq=aln.ptn_freq, kind="nonrev", ntaxa=ntaxa,
                   blen=np.array([(a, b, tree.get_length(a, b)) for a, b in tree.branches()]),
                   adj_pad=np.array([list(x) + [-1] * (3 - len(x)) for x in tree.adj], dtype=np.int32),
                   Q=nm.Q, ceval=nm.ceval, cevec=nm.cevec, cinv_evec=nm.cinv_evec, freqs=nm.freqs, mixw=nm.mixw,
                   rates=rates, props=props)
        cases = [(ntaxa, tree.adj[ntaxa][0]), (0, tree.adj[0][0])] + internal_branches(tree)[:3]
        res = []
        for a, b in cases:
            if tree.is_leaf(a):
                lnl = ref.lnl_at_leaf(a)
            else:
                lnl = ref.lnl_at_branch(leaf_set(tree, b, a), from_taxon=leaf_set(tree, a, b)[0])
            df, ddf = (0.0, 0.0) if a == ntaxa else ref.derv()
            res.append((a, b, lnl, df, ddf))
        out["cases"] = np.array(res)
        out["pattern_lh"] = ref.pattern_lh()
        ref.close()
        save(name, out)


if __name__ == "__main__":
    main()
