"""CPU oracle for the tree-likelihood path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  ``build()`` compiles the C restatement
(phylokernel_oracle.c) and, when /root/reference is present, the reference's own
kernel translation units into oracle/_ref/ (see oracle/Makefile).
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_LIB = os.path.join(HERE, "liboracle.so")
REF_LIB = os.path.join(HERE, "_ref", "libiqtree_ref.so")


def build(with_ref: bool = True, verbose: bool = False) -> None:
    targets = ["liboracle.so"]
    if with_ref and os.path.isdir("/root/reference/tree") and os.path.exists(os.path.join(HERE, "Makefile.ref")):
        targets.append("ref")
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-s", "-C", HERE, "-j8"] + targets, stdout=out)


def oracle_engine(aln, model, rates, props, **kw):
    """An Engine (iqtree2_b200.capi) backed by the C restatement."""
    from iqtree2_b200.capi import engine_for
    if not os.path.exists(ORACLE_LIB):
        build(with_ref=False)
    return engine_for(aln, model, rates, props, lib_path=ORACLE_LIB, prefix="orc_", **kw)


def set_ref_quirks(on: bool) -> None:
    """Toggle the reference's odd-nstates dotProductTriple quirk in the restatement (default on)."""
    import ctypes
    if not os.path.exists(ORACLE_LIB):
        build(with_ref=False)
    ctypes.CDLL(ORACLE_LIB).orc_set_ref_quirks(int(bool(on)))


# ---- the reference's own kernels (oracle/_ref, built by Makefile.ref where /root/reference exists) ----------
LK_SSE2, LK_AVX, LK_AVX_FMA = 2, 7, 8   # enum LikelihoodKernel, utils/tools.h:526-528
LK_CUDA = 10                            # appended by adapter/lk_cuda.patch; only adapter/_build/libiqtree_cuda.so knows it
# the reference's own PhyloTree with the LK_CUDA selector patch and the adapter TU (adapter/Makefile): same harness,
# SSE/AVX/FMA kernels AND the four pointers bound to libb200lk.so
ADAPTER_LIB = os.path.join(os.path.dirname(HERE), "adapter", "_build", "libiqtree_cuda.so")


def adapter_available() -> bool:
    return os.path.exists(ADAPTER_LIB)


def build_adapter(verbose: bool = False) -> None:
    """adapter/_build/libiqtree_cuda.so -- needs /root/reference (this container) and the built libb200lk.so."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-s", "-C", os.path.join(os.path.dirname(HERE), "adapter"), "-j8"], stdout=out)


def ref_available() -> bool:
    return os.path.exists(REF_LIB)


class RefRun:
    """One PhyloTree of the reference with caller-supplied alignment, tree, eigen-system and rates
    (ref_harness.cpp).  kernel: LK_SSE2 (Vec2d), LK_AVX (Vec4d), LK_AVX_FMA (Vec4d + FMA)."""

    def __init__(self, tree, aln, model, rates, props, *, kernel=LK_AVX_FMA, safe=False, threads=1, p_invar=0.0,
                 unobserved=None, fused=False, site_models=None, nonrev=None, mixlen=None, asc_holder=False,
                 robust=None, lib_path=None):
        import ctypes as C
        import numpy as np
        from iqtree2_b200 import workload as W
        if lib_path is None and not ref_available():
            raise RuntimeError("oracle/_ref/libiqtree_ref.so is not built (needs /root/reference)")
        self.C, self.np = C, np
        self.lib = C.CDLL(lib_path or REF_LIB)

        class Problem(C.Structure):
            _fields_ = [(k, C.c_int32) for k in ("ntaxa", "nstates", "ncat", "nmix", "n_state_codes", "kernel",
                                                 "safe", "threads")] + \
                       [("nptn", C.c_int64), ("states", C.c_void_p), ("ptn_freq", C.c_void_p),
                        ("newick", C.c_char_p), ("tipvec", C.c_void_p), ("rates", C.c_void_p),
                        ("props", C.c_void_p), ("pinvar", C.c_double), ("eval", C.c_void_p), ("evec", C.c_void_p),
                        ("inv_evec", C.c_void_p), ("mixw", C.c_void_p), ("freqs", C.c_void_p)]
        names = ["T%d" % i for i in range(aln.ntaxa)]
        old = tree.names
        tree.names = names
        if nonrev is not None and nonrev[0] == 1:
            # rooted: `tree` carries the root as its LAST leaf (id ntaxa of the alignment); the reference reads a newick
            # whose top node has two children and adds that leaf itself (tree/mtree.cpp:739-751)
            tree.names = names + ["__root__"]
            newick = tree.to_newick_rooted(tree.ntaxa - 1, digits=17).encode()
        else:
            newick = tree.to_newick(0, digits=17).encode()
        tree.names = old
        f64 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        scale = 1.0 - p_invar
        self._keep = [np.ascontiguousarray(aln.states, dtype=np.uint8), f64(aln.ptn_freq),
                      f64(W.state_indicators(aln.nstates)), f64(rates), f64(np.asarray(props) * scale),
                      f64(model.eval), f64(model.evec), f64(model.inv_evec), f64(model.mixw), f64(model.freqs)]
        k = self._keep
        ptr = lambda a: a.ctypes.data
        pb = Problem(aln.ntaxa, aln.nstates, len(rates), model.nmix, aln.state_unknown + 1, kernel, int(safe),
                     threads, aln.nptn, ptr(k[0]), ptr(k[1]), newick, ptr(k[2]), ptr(k[3]), ptr(k[4]), p_invar,
                     ptr(k[5]), ptr(k[6]), ptr(k[7]), ptr(k[8]), ptr(k[9]))
        if unobserved is not None:  # +ASC (Lewis): [ntaxa][n_unobserved] state codes
            un = np.ascontiguousarray(unobserved, dtype=np.uint8)
            self.lib.ref_set_next_unobserved(un.shape[1], C.c_void_p(un.ctypes.data), un.shape[0])
        if site_models is not None:  # (eval [nptn][n], evec [nptn][n][n], inv_evec [nptn][n][n]): ModelSet / PMSF
            sm = [np.ascontiguousarray(a, dtype=np.float64) for a in site_models]
            self.lib.ref_set_next_site_models(C.c_int64(aln.nptn), aln.nstates, *[C.c_void_p(a.ctypes.data) for a in sm])
        if nonrev is not None:  # (mode, Q, ceval, cevec, cinv_evec): mode 1 rooted + non-reversible model, 2 kernel_nonrev
            cx = [np.ascontiguousarray(a, dtype=np.complex128) for a in nonrev[2:5]]
            q = f64(nonrev[1])
            self.lib.ref_set_next_nonrev(int(nonrev[0]), aln.nstates, model.nmix, C.c_void_p(q.ctypes.data),
                                         *[C.c_void_p(a.ctypes.data) for a in cx])
        if mixlen is not None:  # heterotachy: class c of every branch has length (newick length) * mixlen[c]
            mm = f64(mixlen)
            self.lib.ref_set_next_mixlen(len(mm), C.c_void_p(mm.ctypes.data))
        if robust is not None:  # (keep, median): params->robust_phy_keep / robust_median
            self.lib.ref_set_next_robust(C.c_double(robust[0]), int(robust[1]))
        if asc_holder:  # the unobserved patterns are Holder's (ASC_VARIANT_MISSING)
            self.lib.ref_set_next_asc_holder(1)
        if fused:  # model_factory->fused_mix_rate: category c = mixture class c at rate category c
            self.lib.ref_set_next_fused(1)
        self.lib.ref_create.restype = C.c_void_p
        self.h = C.c_void_p(self.lib.ref_create(C.byref(pb)))
        self.nptn, self.ncm, self.n = aln.nptn, (len(rates) if fused else len(rates) * model.nmix), aln.nstates
        self.tree = tree
        for fn in ("ref_lnl_at_leaf", "ref_lnl_again", "ref_from_buffer", "ref_optimize_current_branch",
                   "ref_optimize_all_branches", "ref_compute_likelihood", "ref_branch_length", "ref_lnl_at_branch"):
            getattr(self.lib, fn).restype = C.c_double
        self.safe_numeric = bool(self.lib.ref_safe_numeric(self.h))
        self.vector_size = self.lib.ref_vector_size(self.h)

    def optimize_all_branches(self, iterations=100, tolerance=0.001):
        return self.lib.ref_optimize_all_branches(self.h, iterations, self.C.c_double(tolerance))

    def compute_likelihood(self):
        return self.lib.ref_compute_likelihood(self.h)

    def branch_length(self, leaves, from_taxon=0):
        arr = self.np.ascontiguousarray(sorted(leaves), dtype=self.np.int32)
        return self.lib.ref_branch_length(self.h, self.C.c_void_p(arr.ctypes.data), len(arr), from_taxon)

    def nni(self, leaves, from_taxon=0, nni5=False):
        """getBestNNIForBran on the internal branch behind `leaves`: -> (winner, [lnL NNI 1, lnL NNI 2], [central
        lengths], lnL before, lnL after restoring)."""
        np, C = self.np, self.C
        arr = np.ascontiguousarray(sorted(leaves), dtype=np.int32)
        ll, ln = np.zeros(2), np.zeros(2)
        before, after = C.c_double(), C.c_double()
        w = self.lib.ref_nni(self.h, C.c_void_p(arr.ctypes.data), len(arr), from_taxon, int(nni5), C.c_void_p(ll.ctypes.data),
                             C.c_void_p(ln.ctypes.data), C.byref(before), C.byref(after))
        if w < 0:
            raise RuntimeError("ref_nni failed (%d)" % w)
        return w, ll, ln, before.value, after.value

    def lnl_at_leaf(self, taxon):
        return self.lib.ref_lnl_at_leaf(self.h, taxon)

    def lnl_at_branch(self, leaves, from_taxon=0):
        """lnL at the internal branch whose far side (seen from leaf ``from_taxon``) has exactly ``leaves``."""
        arr = self.np.ascontiguousarray(sorted(leaves), dtype=self.np.int32)
        return self.lib.ref_lnl_at_branch(self.h, self.C.c_void_p(arr.ctypes.data), len(arr), from_taxon)

    def derv(self, again=False):
        C = self.C
        df, ddf = C.c_double(), C.c_double()
        (self.lib.ref_derv_again if again else self.lib.ref_derv)(self.h, C.byref(df), C.byref(ddf))
        return df.value, ddf.value

    def set_current_length(self, t):
        self.lib.ref_set_current_length(self.h, self.C.c_double(t))

    # heterotachy
    def set_current_mixlen(self, lens):
        a = self.np.ascontiguousarray(lens, dtype=self.np.float64)
        self.lib.ref_set_current_mixlen(self.h, self.C.c_void_p(a.ctypes.data), len(a))

    def derv_mixlen(self, nclass, again=False):
        np, C = self.np, self.C
        df, ddf, lnl = np.zeros(nclass), np.zeros((nclass, nclass)), C.c_double()
        self.lib.ref_derv_mixlen(self.h, C.c_void_p(df.ctypes.data), C.c_void_p(ddf.ctypes.data), C.byref(lnl),
                                 int(again))
        return df, ddf, lnl.value

    def derv_mixlen_class(self, cls, length, again=False):
        C = self.C
        df, ddf = C.c_double(), C.c_double()
        self.lib.ref_derv_mixlen_class(self.h, int(cls), C.c_double(length), C.byref(df), C.byref(ddf), int(again))
        return df.value, ddf.value

    def set_ptn_freq(self, freq):
        a = self.np.ascontiguousarray(freq, dtype=self.np.float64)
        self.lib.ref_set_ptn_freq(self.h, self.C.c_void_p(a.ctypes.data))

    def from_buffer(self):
        return self.lib.ref_from_buffer(self.h)

    def optimize_current_branch(self, max_steps=100):
        return self.lib.ref_optimize_current_branch(self.h, max_steps)

    def pattern_lh(self):
        out = self.np.empty(self.nptn)
        self.lib.ref_pattern_lh(self.h, out.ctypes.data_as(self.C.c_void_p))
        return out

    def pattern_lh_cat(self):
        out = self.np.empty((self.nptn, self.ncm))
        self.lib.ref_pattern_lh_cat(self.h, out.ctypes.data_as(self.C.c_void_p))
        return out

    def fetch_branch(self, leaves, from_taxon):
        """(partial [nptn][ncm][n], scale) of the directed branch whose subtree has exactly `leaves`."""
        np, C = self.np, self.C
        ls = np.ascontiguousarray(sorted(leaves), dtype=np.int32)
        part = np.empty((self.nptn, self.ncm, self.n))
        sc = np.empty((self.nptn, self.ncm) if self.safe_numeric else (self.nptn,), dtype=np.uint8)
        rc = self.lib.ref_fetch_branch(self.h, ls.ctypes.data_as(C.c_void_p), len(ls), from_taxon,
                                       part.ctypes.data_as(C.c_void_p), sc.ctypes.data_as(C.c_void_p))
        if rc:
            raise RuntimeError("ref_fetch_branch failed (%d)" % rc)
        return part, sc

    def close(self):
        if self.h:
            self.lib.ref_destroy(self.h)
            self.h = None
