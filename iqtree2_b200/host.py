"""ctypes driver of the C++ host layer (iqtree2_b200/host/ -> libb200host.so): the reference's PhyloTree /
PhyloSuperTree dispatch API over LK_CUDA.  Used by tests/ and bench.py; the host layer itself is C++."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb200host.so")
SEQ_DNA, SEQ_PROTEIN, SEQ_MORPH = 0, 1, 2
LK_AVX_FMA, LK_CUDA = 4, 6
_lib = None


class HostError(RuntimeError):
    pass


def _load(path):
    if not os.path.exists(path):
        raise HostError("%s is missing: build it first (python -m iqtree2_b200.build_host)" % path)
    L = C.CDLL(path)
    L.b200host_last_error.restype = C.c_char_p
    L.b200host_alignment_npattern.restype = C.c_int64
    L.b200host_alignment_name.restype = C.c_char_p
    L.b200host_tree_kernel_calls.restype = C.c_int64
    L.b200host_tree_gpu_launches.restype = C.c_int64
    return L


def lib():
    """The host library in use: the product build, unless a test has switched to the testing build."""
    global _lib
    if _lib is None:
        _lib = _load(LIB_PATH)
    return _lib


def _ck(rc):
    if rc != 0:
        raise HostError((lib().b200host_last_error() or b"").decode())


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


_product_lib = None
_testing_lib = None


def set_kernel_library(path=None, prefix=None):
    """TEST HOOK.  Bind the host layer to another library exporting the C ABI (tests pass the CPU oracle).  The
    product build (libb200host.so) has no such switch: this loads the -DB200_TESTING build of the same sources
    (libb200host_testing.so) and makes it the library `lib()` returns; None restores the product library.  Objects
    created under one library must be released before switching."""
    global _lib, _product_lib, _testing_lib
    if path is None and prefix is None:
        if _testing_lib is not None:
            _ck_on(_testing_lib, _testing_lib.b200host_set_kernel_library(b"", b""))
        if _product_lib is None:
            _product_lib = _load(LIB_PATH)
        _lib = _product_lib
        return
    if _testing_lib is None:
        from . import build_host
        _testing_lib = _load(build_host.build(testing=True))
    _lib = _testing_lib
    _ck(_lib.b200host_set_kernel_library((path or "").encode(), (prefix or "").encode()))


def _ck_on(L, rc):
    if rc != 0:
        raise HostError((L.b200host_last_error() or b"").decode())


def _seq_type(nstates):
    return SEQ_DNA if nstates == 4 else (SEQ_PROTEIN if nstates == 20 else SEQ_MORPH)


class HostAlignment:
    def __init__(self, handle):
        self.h = handle

    @classmethod
    def from_states(cls, codes, nstates, compress=True):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        h = C.c_void_p()
        _ck(lib().b200host_alignment_from_states(C.byref(h), codes.shape[0], C.c_int64(codes.shape[1]),
                                                 codes.ctypes.data_as(C.c_void_p), _seq_type(nstates), nstates,
                                                 int(compress)))
        return cls(h)

    @classmethod
    def from_phylip(cls, path, seq_type=SEQ_DNA):
        h = C.c_void_p()
        _ck(lib().b200host_alignment_from_phylip(C.byref(h), path.encode(), seq_type))
        return cls(h)

    @property
    def nptn(self):
        return lib().b200host_alignment_npattern(self.h)

    @property
    def ntaxa(self):
        return lib().b200host_alignment_ntaxa(self.h)

    def names(self):
        return [lib().b200host_alignment_name(self.h, i).decode() for i in range(self.ntaxa)]

    def patterns(self):
        st = np.empty((self.ntaxa, self.nptn), dtype=np.uint8)
        fr = np.empty(self.nptn)
        cc = np.empty(self.nptn, dtype=np.int32)
        lib().b200host_alignment_get(self.h, st.ctypes.data_as(C.c_void_p), _dp(fr), cc.ctypes.data_as(C.c_void_p))
        return st, fr, cc


class HostModel:
    def __init__(self, handle, n, nmix):
        self.h, self.n, self.nmix = handle, n, nmix

    @classmethod
    def gtr(cls, rates, freqs):
        h = C.c_void_p()
        r, f = np.asarray(rates, dtype=np.float64), np.asarray(freqs, dtype=np.float64)
        _ck(lib().b200host_model_gtr(C.byref(h), _dp(r), _dp(f)))
        return cls(h, 4, 1)

    @classmethod
    def named(cls, name):
        """The protein models BASELINE names, from the data compiled into the host layer: "LG", "LG+C60" (and "JC")."""
        h = C.c_void_p()
        _ck(lib().b200host_model_named(C.byref(h), name.encode()))
        n, nmix = {"LG": (20, 1), "LG+C60": (20, 60), "JC": (4, 1)}[name]
        return cls(h, n, nmix)

    @classmethod
    def reversible(cls, exch, freqs):
        e, f = np.ascontiguousarray(exch, dtype=np.float64), np.asarray(freqs, dtype=np.float64)
        h = C.c_void_p()
        _ck(lib().b200host_model_reversible(C.byref(h), len(f), _dp(e), _dp(f)))
        return cls(h, len(f), 1)

    @classmethod
    def mixture(cls, classes, weights):
        arr = (C.c_void_p * len(classes))(*[c.h for c in classes])
        w = np.asarray(weights, dtype=np.float64)
        h = C.c_void_p()
        _ck(lib().b200host_model_mixture(C.byref(h), arr, _dp(w), len(classes)))
        return cls(h, classes[0].n, len(classes))

    @classmethod
    def from_eigen(cls, model):
        """model: workload.EigenModel"""
        a = [np.ascontiguousarray(x, dtype=np.float64) for x in
             (model.eval, model.evec, model.inv_evec, model.freqs, model.mixw)]
        h = C.c_void_p()
        _ck(lib().b200host_model_from_eigen(C.byref(h), model.nstates, model.nmix, *[_dp(x) for x in a]))
        return cls(h, model.nstates, model.nmix)

    def eigen(self):
        n, m = self.n, self.nmix
        ev, U, Ui, fr, w = np.empty((m, n)), np.empty((m, n, n)), np.empty((m, n, n)), np.empty((m, n)), np.empty(m)
        lib().b200host_model_get(self.h, _dp(ev), _dp(U), _dp(Ui), _dp(fr), _dp(w))
        return ev, U, Ui, fr, w


class HostRate:
    def __init__(self, handle, ncat):
        self.h, self.ncat = handle, ncat

    @classmethod
    def gamma(cls, alpha, ncat=4, p_invar=0.0):
        h = C.c_void_p()
        _ck(lib().b200host_rate_gamma(C.byref(h), C.c_double(alpha), ncat, C.c_double(p_invar)))
        return cls(h, ncat)

    @classmethod
    def categories(cls, rates, props, p_invar=0.0):
        r, p = np.asarray(rates, dtype=np.float64), np.asarray(props, dtype=np.float64)
        h = C.c_void_p()
        _ck(lib().b200host_rate_categories(C.byref(h), _dp(r), _dp(p), len(r), C.c_double(p_invar)))
        return cls(h, len(r))

    def get(self):
        r, p = np.empty(self.ncat), np.empty(self.ncat)
        lib().b200host_rate_get(self.h, _dp(r), _dp(p))
        return r, p


class HostTree:
    """b200::PhyloTree set up in the reference's order (readTree, setAlignment, setModel, setLikelihoodKernel
    (LK_CUDA), initializeAllPartialLh)."""

    def __init__(self, newick, aln, model, rate, safe=False, device=0, device_newton=True, asc=False, mem_save=None,
                 asc_holder=False, robust=None, mixlen=None, devices=None):
        """mem_save: None = one slot per internal node; a number = -mem (LM_MEM_SAVE) with that Params::max_mem_size
        (0: the minimum log2(ntaxa)+1 slots; <= 1: that fraction of ntaxa-2).  asc_holder: +ASC with missing data.
        robust: (keep, median) = Params::robust_phy_keep / robust_median.  mixlen: class multipliers -> a
        b200::PhyloTreeMixlen (heterotachy) whose class c of every branch starts at length x mixlen[c]."""
        self.h = C.c_void_p()
        self._keep = (aln, model, rate)
        if mem_save is not None:
            lib().b200host_set_mem_fraction(C.c_double(float(mem_save)))
        if robust is not None:
            lib().b200host_set_robust(C.c_double(float(robust[0])), int(bool(robust[1])))
        if devices is not None:  # Params::devices: the alignment split by site pattern over these GPUs
            arr = (C.c_int32 * len(devices))(*devices)
            lib().b200host_set_next_devices(arr, len(devices))
        self.mixlen = 0
        if mixlen is not None:
            mm = np.ascontiguousarray(mixlen, dtype=np.float64)
            lib().b200host_set_next_mixlen(_dp(mm), len(mm))
            self.mixlen = len(mm)
        _ck(lib().b200host_tree_new(C.byref(self.h), newick.encode(), aln.h, model.h, rate.h,
                                    int(bool(safe)) | (2 if asc else 0) | (4 if mem_save is not None else 0) |
                                    (8 if asc_holder else 0), device, int(device_newton)))
        self.aln = aln

    def close(self):
        if self.h:
            lib().b200host_tree_free(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def max_lh_slots(self):
        return lib().b200host_tree_max_lh_slots(self.h)

    @property
    def nbranches(self):
        return lib().b200host_tree_nbranches(self.h)

    def set_branch_bounds(self, lo, hi):
        lib().b200host_tree_set_min_branch(self.h, C.c_double(lo), C.c_double(hi))

    def compute_likelihood(self, want_pattern_lh=False):
        out = C.c_double()
        plh = np.empty(self.aln.nptn) if want_pattern_lh else None
        _ck(lib().b200host_tree_compute_likelihood(self.h, _dp(plh) if want_pattern_lh else None, C.byref(out)))
        return (out.value, plh) if want_pattern_lh else out.value

    def clear_partials(self):
        _ck(lib().b200host_tree_clear_partials(self.h))

    def branch_lnl(self, bid):
        out = C.c_double()
        _ck(lib().b200host_tree_branch_lnl(self.h, bid, C.byref(out)))
        return out.value

    def derv(self, bid, length):
        df, ddf = C.c_double(), C.c_double()
        _ck(lib().b200host_tree_derv(self.h, bid, C.c_double(length), C.byref(df), C.byref(ddf)))
        return df.value, ddf.value

    def optimize_one_branch(self, bid, max_steps=100):
        out = C.c_double()
        _ck(lib().b200host_tree_optimize_one_branch(self.h, bid, max_steps, C.byref(out)))
        return out.value

    def pattern_lh_cat(self, wsl=2):
        """computePatternLhCat(wsl): -> (lnL, [nptn][columns]); wsl 2 = WSL_RATECAT, 3 = WSL_MIXTURE, 4 = both."""
        ncols, lnl = C.c_int32(), C.c_double()
        _ck(lib().b200host_tree_pattern_lh_cat(self.h, wsl, None, C.byref(ncols), C.byref(lnl)))
        out = np.empty((self.aln.nptn, ncols.value))
        _ck(lib().b200host_tree_pattern_lh_cat(self.h, wsl, _dp(out), C.byref(ncols), C.byref(lnl)))
        return lnl.value, out

    def derv_multi(self, bid, lens):
        """heterotachy: gradient [mixlen], Hessian [mixlen][mixlen] and lnL at class lengths `lens` of branch bid."""
        a = np.ascontiguousarray(lens, dtype=np.float64)
        df, ddf, lnl = np.zeros(self.mixlen), np.zeros((self.mixlen, self.mixlen)), C.c_double()
        _ck(lib().b200host_tree_derv_multi(self.h, bid, _dp(a), _dp(df), _dp(ddf), C.byref(lnl)))
        return df, ddf, lnl.value

    def mix_lengths(self):
        v = np.empty((self.nbranches, self.mixlen))
        lib().b200host_tree_get_mix_lengths(self.h, _dp(v))
        return v

    def from_buffer(self):
        out = C.c_double()
        _ck(lib().b200host_tree_from_buffer(self.h, C.byref(out)))
        return out.value

    def optimize_all_branches(self, iterations=100, tolerance=0.001):
        out = C.c_double()
        _ck(lib().b200host_tree_optimize_all_branches(self.h, iterations, C.c_double(tolerance), C.byref(out)))
        return out.value

    def lengths(self):
        v = np.empty(self.nbranches)
        lib().b200host_tree_get_lengths(self.h, _dp(v))
        return v

    def set_lengths(self, v):
        v = np.ascontiguousarray(v, dtype=np.float64)
        _ck(lib().b200host_tree_set_lengths(self.h, _dp(v)))

    def branch_split(self, bid):
        rows = np.empty(self.aln.ntaxa, dtype=np.int32)
        n = C.c_int32()
        _ck(lib().b200host_tree_branch_split(self.h, bid, rows.ctypes.data_as(C.c_void_p), C.byref(n)))
        return frozenset(int(x) for x in rows[:n.value])

    def newick(self):
        buf = C.create_string_buffer(1 << 22)
        _ck(lib().b200host_tree_newick(self.h, buf, len(buf)))
        return buf.value.decode()

    def set_device_sweep(self, on):
        """Params::device_sweep: optimizeAllBranches keeps the lengths on the device for a whole sweep (default on)."""
        lib().b200host_tree_set_device_sweep(self.h, int(bool(on)))

    def device_sweeps(self):
        f = lib().b200host_tree_device_sweeps
        f.restype = C.c_int64
        return f(self.h)

    def kernel_calls(self):
        return lib().b200host_tree_kernel_calls(self.h)

    def gpu_launches(self):
        """Kernels launched by this tree's device context(s) since creation."""
        return lib().b200host_tree_gpu_launches(self.h)

    def safe_numeric(self):
        return bool(lib().b200host_tree_safe_numeric(self.h))

    def set_kernel(self, lk):
        _ck(lib().b200host_tree_set_kernel(self.h, lk))


class HostSuperTree(HostTree):
    """b200::PhyloSuperTree: partitions over one topology (edge-linked proportional, or unlinked)."""

    def __init__(self, newick, edge_linked=True, device=0):
        self.h = C.c_void_p()
        self._keep = []
        self.nparts = 0
        _ck(lib().b200host_super_new(C.byref(self.h), newick.encode(), int(edge_linked), device))

    def add_partition(self, name, aln, model, rate, part_rate=1.0):
        self._keep.append((aln, model, rate))
        _ck(lib().b200host_super_add(self.h, name.encode(), aln.h, model.h, rate.h, C.c_double(part_rate)))
        self.nparts += 1

    def set_batching(self, on):
        """Batched launches over partitions of equal dimensions (default on); call before initialize()."""
        lib().b200host_super_set_batching(self.h, int(on))

    def batch_info(self):
        """-> (number of batches, partitions inside batches, batched calls made so far)"""
        nb, npart, calls = C.c_int32(), C.c_int32(), C.c_int64()
        lib().b200host_super_batch_info(self.h, C.byref(nb), C.byref(npart), C.byref(calls))
        return nb.value, npart.value, calls.value

    def initialize(self):
        _ck(lib().b200host_super_init(self.h))

    def compute_likelihood(self, want_pattern_lh=False):
        out = C.c_double()
        _ck(lib().b200host_tree_compute_likelihood(self.h, None, C.byref(out)))
        return out.value

    def part_scores(self):
        v = np.empty(self.nparts)
        lib().b200host_super_part_scores(self.h, _dp(v))
        return v

    def part_order(self):
        v = np.empty(self.nparts, dtype=np.int32)
        lib().b200host_super_part_order(self.h, v.ctypes.data_as(C.c_void_p))
        return v
