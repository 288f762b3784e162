"""ctypes binding of the C ABI in include/b200lk.h.

The same binding class drives any library that exports the ABI under a symbol
prefix: the product (``libb200lk.so``, prefix ``b200lk_``) and -- from tests
only -- the CPU oracle (prefix ``orc_``).  There is no fallback: loading the
product library fails loudly when it has not been built or cannot be loaded.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# B200LK_LIB: A/B hook for kernel experiments (another build of the SAME library); the default is the in-tree build
LIB_PATH = os.environ.get("B200LK_LIB") or os.path.join(HERE, "libb200lk.so")

LEAF, INTERNAL = 0, 1
ROOT, TOWARD_ROOT = 2, 0x100          # non-reversible contexts: the root leaf; child->direction == TOWARD_ROOT
MODEL_REV, MODEL_SITE, MODEL_NONREV = 0, 1, 2   # b200lk_dims.site_specific
WARN_LNL_NONFINITE, WARN_DERV_NONFINITE, INFO_DERV_NOT_AS_REFERENCE = 1, 2, 4


class Dims(C.Structure):
    _fields_ = [("ntaxa", C.c_int32), ("nstates", C.c_int32), ("ncat", C.c_int32), ("nmix", C.c_int32),
                ("fused_mix_rate", C.c_int32), ("safe_numeric", C.c_int32), ("n_state_codes", C.c_int32),
                ("nslots", C.c_int32), ("nptn", C.c_int64), ("n_unobserved", C.c_int64),
                ("device", C.c_int32), ("site_specific", C.c_int32)]


class Node(C.Structure):
    _fields_ = [("out_slot", C.c_int32), ("child_begin", C.c_int32), ("nchild", C.c_int32),
                ("reserved", C.c_int32)]


class Child(C.Structure):
    _fields_ = [("kind", C.c_int32), ("id", C.c_int32), ("length", C.c_double)]


class End(C.Structure):
    _fields_ = [("kind", C.c_int32), ("id", C.c_int32)]


class B200LKError(RuntimeError):
    pass


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _u8p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def pack_traversal(entries):
    """entries: [(node, dad, out_slot, [(kind, id, length), ...]), ...] (workload.traversal)
    -> (Node array, nnodes, Child array, nchildren)."""
    nch = sum(len(e[3]) for e in entries)
    nodes = (Node * max(1, len(entries)))()
    kids = (Child * max(1, nch))()
    j = 0
    for i, e in enumerate(entries):
        nodes[i] = Node(e[2], j, len(e[3]), 0)
        for kind, cid, ln in e[3]:
            kids[j] = Child(kind, cid, ln)
            j += 1
    return nodes, len(entries), kids, nch


class Engine:
    """One likelihood context.  Methods mirror include/b200lk.h one to one."""

    def __init__(self, *, ntaxa, nstates, ncat, nmix=1, nptn, nslots, n_state_codes, safe_numeric=False,
                 fused_mix_rate=False, n_unobserved=0, device=0, lib_path=LIB_PATH, prefix="b200lk_",
                 site_specific=False, nonrev=False, devices=None):
        """devices: a list of CUDA ordinals -> b200lk_create_sharded (the problem split by site pattern over them, all
        driven from this process); None -> b200lk_create on `device`."""
        if not os.path.exists(lib_path):
            raise B200LKError(
                "%s is missing: build it first (python -c 'import __graft_entry__ as g; g.build()'). "
                "There is no CPU fallback on this path." % lib_path)
        self.lib = C.CDLL(lib_path)
        self.prefix = prefix
        self.dims = Dims(ntaxa, nstates, ncat, nmix, int(fused_mix_rate), int(safe_numeric), n_state_codes,
                         nslots, nptn, n_unobserved, device, MODEL_NONREV if nonrev else int(site_specific))
        self.ncat_mix = ncat if fused_mix_rate else ncat * nmix
        self.block = nstates * self.ncat_mix
        self.ctx = C.c_void_p()
        if devices is not None:
            arr = (C.c_int32 * len(devices))(*devices)
            self._check(self._f("create_sharded")(C.byref(self.ctx), C.byref(self.dims), arr, len(devices)))
        else:
            self._check(self._f("create")(C.byref(self.ctx), C.byref(self.dims)))
        self._keep = []

    # -- plumbing ------------------------------------------------------------
    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    def _check(self, rc):
        if rc < 0:
            msg = ""
            if self.prefix == "b200lk_" and self.ctx:
                self.lib.b200lk_last_error.restype = C.c_char_p
                msg = (self.lib.b200lk_last_error(self.ctx) or b"").decode()
            raise B200LKError("%s call failed with status %d %s" % (self.prefix, rc, msg))
        return rc

    def close(self):
        if self.ctx:
            self._f("destroy")(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- inputs --------------------------------------------------------------
    def upload_tips(self, states, ptn_freq, unobserved=None):
        states = np.ascontiguousarray(states, dtype=np.uint8)
        ptn_freq = np.ascontiguousarray(ptn_freq, dtype=np.float64)
        assert states.shape == (self.dims.ntaxa, self.dims.nptn)
        un = None
        if unobserved is not None:
            unobserved = np.ascontiguousarray(unobserved, dtype=np.uint8)
            un = _u8p(unobserved)
        self._check(self._f("upload_tips")(self.ctx, _u8p(states), un, _dp(ptn_freq)))

    def set_ptn_invar(self, ptn_invar):
        a = np.ascontiguousarray(ptn_invar, dtype=np.float64)
        assert a.size == self.dims.nptn + self.dims.n_unobserved
        self._check(self._f("set_ptn_invar")(self.ctx, _dp(a)))

    def set_model(self, eval_, evec, inv_evec, rates, props, mixw, tip_partial_lh):
        arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in
                (eval_, evec, inv_evec, rates, props, mixw, tip_partial_lh)]
        n, nmix = self.dims.nstates, self.dims.nmix
        assert arrs[0].size == nmix * n and arrs[1].size == nmix * n * n and arrs[2].size == nmix * n * n
        assert arrs[3].size == self.dims.ncat and arrs[4].size == self.dims.ncat and arrs[5].size == nmix
        assert arrs[6].size == self.dims.n_state_codes * nmix * n
        self._check(self._f("set_model")(self.ctx, *[_dp(a) for a in arrs]))

    # -- kernels -------------------------------------------------------------
    def set_model_rates(self, exch, freqs, rates, props, mixw, tipvec, total_num_subst=None):
        """Model from its free parameters; the eigen-decomposition runs on the device (b200lk_set_model_rates).
        exch [nmix][n(n-1)/2] upper triangle, freqs [nmix][n], tipvec [n_state_codes][n]."""
        arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (exch, freqs, tipvec, rates, props, mixw)]
        tot = None if total_num_subst is None else np.ascontiguousarray(total_num_subst, dtype=np.float64)
        f = self._f("set_model_rates")
        f.argtypes = [C.c_void_p] + [C.POINTER(C.c_double)] * 7
        self._check(f(self.ctx, _dp(arrs[0]), _dp(arrs[1]), None if tot is None else _dp(tot), _dp(arrs[2]),
                      _dp(arrs[3]), _dp(arrs[4]), _dp(arrs[5])))

    def set_site_models(self, eval_, evec, inv_evec, rates, props, tipvec):
        """Site-specific models (PMSF / ModelSet): eval [nptn][n], evec / inv_evec [nptn][n][n], tipvec [codes][n]."""
        arrs = [np.ascontiguousarray(a, dtype=np.float64) for a in (eval_, evec, inv_evec, rates, props, tipvec)]
        f = self._f("set_site_models")
        f.argtypes = [C.c_void_p] + [C.POINTER(C.c_double)] * 6
        self._check(f(self.ctx, *[_dp(a) for a in arrs]))

    def set_model_nonrev(self, rate_matrix, ceval, cevec, cinv_evec, freqs, tipvec, rates, props, mixw, rooted):
        """Non-reversible kernel family: Q [nmix][n][n], complex eigen-system of Q, state frequencies."""
        n, nmix = self.dims.nstates, self.dims.nmix
        cx = [np.ascontiguousarray(a, dtype=np.complex128).view(np.float64) for a in (ceval, cevec, cinv_evec)]
        re = [np.ascontiguousarray(a, dtype=np.float64) for a in (rate_matrix, freqs, tipvec, rates, props, mixw)]
        assert re[0].size == nmix * n * n and cx[0].size == 2 * nmix * n and cx[1].size == 2 * nmix * n * n
        assert cx[2].size == 2 * nmix * n * n and re[1].size == nmix * n and re[2].size == self.dims.n_state_codes * n
        self._check(self._f("set_model_nonrev")(self.ctx, _dp(re[0]), _dp(cx[0]), _dp(cx[1]), _dp(cx[2]), _dp(re[1]),
                                                _dp(re[2]), _dp(re[3]), _dp(re[4]), _dp(re[5]), int(bool(rooted))))

    def fetch_model(self):
        n, nmix = self.dims.nstates, self.dims.nmix
        ev, U, Ui = np.zeros((nmix, n)), np.zeros((nmix, n, n)), np.zeros((nmix, n, n))
        f = self._f("fetch_model")
        f.argtypes = [C.c_void_p] + [C.POINTER(C.c_double)] * 3
        self._check(f(self.ctx, _dp(ev), _dp(U), _dp(Ui)))
        return ev, U, Ui

    def update_partials(self, entries):
        nodes, nn, kids, nk = pack_traversal(entries)
        return self._check(self._f("update_partials")(self.ctx, nodes, nn, kids, nk))

    def branch_lnl(self, dad, node, length):
        out = C.c_double()
        f = self._f("branch_lnl")
        f.argtypes = [C.c_void_p, End, End, C.c_double, C.POINTER(C.c_double)]
        self.last_status = self._check(f(self.ctx, End(*dad), End(*node), length, C.byref(out)))
        return out.value

    def build_theta(self, dad, node):
        f = self._f("build_theta")
        f.argtypes = [C.c_void_p, End, End]
        return self._check(f(self.ctx, End(*dad), End(*node)))

    def derv(self, length):
        df, ddf = C.c_double(), C.c_double()
        f = self._f("derv")
        f.argtypes = [C.c_void_p, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        self.last_status = self._check(f(self.ctx, length, C.byref(df), C.byref(ddf)))
        return df.value, ddf.value

    def lnl_from_buffer(self, length):
        out = C.c_double()
        f = self._f("lnl_from_buffer")
        f.argtypes = [C.c_void_p, C.c_double, C.POINTER(C.c_double)]
        self.last_status = self._check(f(self.ctx, length, C.byref(out)))
        return out.value

    def lnl_from_buffer_noalloc(self, length):
        """b200lk_lnl_from_buffer_noalloc: the same result from the library's static device scratch (no allocation)."""
        out = C.c_double()
        f = self._f("lnl_from_buffer_noalloc")
        f.argtypes = [C.c_void_p, C.c_double, C.POINTER(C.c_double)]
        self.last_status = self._check(f(self.ctx, length, C.byref(out)))
        return out.value

    # -- +ASC flavour, heterotachy ---------------------------------------------
    def set_asc(self, holder):
        """0 / False: Lewis (default with unobserved patterns); 1 / True: Holder's correction for missing data
        (n_unobserved = nstates * nptn, state-major)."""
        return self._check(self._f("set_asc")(self.ctx, C.c_int32(int(holder))))

    def set_robust(self, keep=1.0, median=False):
        """params->robust_phy_keep / robust_median for branch_lnl (tree/phylokernelnew.h:3096-3135)."""
        f = self._f("set_robust")
        f.argtypes = [C.c_void_p, C.c_double, C.c_int32]
        return self._check(f(self.ctx, keep, int(median)))

    def set_ptn_freq(self, ptn_freq):
        a = np.ascontiguousarray(ptn_freq, dtype=np.float64)
        assert a.size == self.dims.nptn
        return self._check(self._f("set_ptn_freq")(self.ctx, _dp(a)))

    def update_partials_mixlen(self, entries, child_len):
        """Heterotachy: child_len [nchildren][ncat], row j = class lengths of the branch to child j of the packed list."""
        nodes, nn, kids, nk = pack_traversal(entries)
        a = np.ascontiguousarray(child_len, dtype=np.float64)
        assert a.shape == (nk, self.dims.ncat)
        f = self._f("update_partials_mixlen")
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.POINTER(C.c_double)]
        return self._check(f(self.ctx, nodes, nn, kids, nk, _dp(a)))

    def branch_lnl_mixlen(self, dad, node, lens):
        a = np.ascontiguousarray(lens, dtype=np.float64)
        assert a.size == self.dims.ncat
        out = C.c_double()
        f = self._f("branch_lnl_mixlen")
        f.argtypes = [C.c_void_p, End, End, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        self.last_status = self._check(f(self.ctx, End(*dad), End(*node), _dp(a), C.byref(out)))
        return out.value

    def lnl_from_buffer_mixlen(self, lens):
        a = np.ascontiguousarray(lens, dtype=np.float64)
        out = C.c_double()
        f = self._f("lnl_from_buffer_mixlen")
        f.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        self.last_status = self._check(f(self.ctx, _dp(a), C.byref(out)))
        return out.value

    def derv_mixlen(self, lens):
        """-> (df [ncat], ddf [ncat][ncat], lnL): computeLikelihoodDerv under isMixlen()."""
        k = self.dims.ncat
        a = np.ascontiguousarray(lens, dtype=np.float64)
        df, ddf, lnl = np.zeros(k), np.zeros((k, k)), C.c_double()
        f = self._f("derv_mixlen")
        f.argtypes = [C.c_void_p] + [C.POINTER(C.c_double)] * 4
        self.last_status = self._check(f(self.ctx, _dp(a), _dp(df), _dp(ddf), C.byref(lnl)))
        return df, ddf, lnl.value

    def derv_mixlen_class(self, cls, length):
        """-> (df, ddf) of heterotachy class `cls` alone (computeLikelihoodDervMixlen)."""
        df, ddf = C.c_double(), C.c_double()
        f = self._f("derv_mixlen_class")
        f.argtypes = [C.c_void_p, C.c_int32, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        self.last_status = self._check(f(self.ctx, cls, length, C.byref(df), C.byref(ddf)))
        return df.value, ddf.value

    def optimize_branch(self, dad, node, guess, min_len=1e-6, max_len=10.0, xacc=None, max_steps=100):
        """b200lk_optimize_branch: theta + the whole safeguarded Newton-Raphson iteration on the device.
        Returns (optimised length, d2l, derivative evaluations)."""

        class Newton(C.Structure):
            _fields_ = [("min_len", C.c_double), ("max_len", C.c_double), ("xacc", C.c_double),
                        ("guess", C.c_double), ("max_steps", C.c_int32), ("reserved", C.c_int32)]
        nr = Newton(min_len, max_len, min_len if xacc is None else xacc, guess, max_steps, 0)
        optx, d2l, steps = C.c_double(), C.c_double(), C.c_int32()
        f = self._f("optimize_branch")
        f.argtypes = [C.c_void_p, End, End, C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double),
                      C.POINTER(C.c_int32)]
        self.last_status = self._check(f(self.ctx, End(*dad), End(*node), C.byref(nr), C.byref(optx), C.byref(d2l),
                                         C.byref(steps)))
        return optx.value, d2l.value, steps.value

    # -- branch lengths resident on the device --------------------------------------
    def set_lengths(self, lengths):
        a = np.ascontiguousarray(lengths, dtype=np.float64)
        self._nlens = a.size
        self._check(self._f("set_lengths")(self.ctx, _dp(a), C.c_int32(a.size)))

    def get_lengths(self):
        """-> (lengths [n], aux [n][2]: derivative evaluations and status of each slot's last optimisation)"""
        n = self._nlens
        ln, aux = np.empty(n), np.empty((n, 2))
        self._check(self._f("get_lengths")(self.ctx, _dp(ln), _dp(aux), C.c_int32(n)))
        return ln, aux

    def update_partials_slots(self, entries, slots):
        """update_partials with the children's lengths taken from the device table: slots[j] >= 0 names the slot of child j
        of the packed list, -1 keeps the length in the entry."""
        nodes, nn, kids, nk = pack_traversal(entries)
        sl = np.ascontiguousarray(slots, dtype=np.int32)
        assert sl.size == nk
        f = self._f("update_partials_slots")
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p]
        return self._check(f(self.ctx, nodes, nn, kids, nk, sl.ctypes.data_as(C.c_void_p)))

    def optimize_branch_slot(self, dad, node, slot, min_len=1e-6, max_len=10.0, xacc=None, max_steps=100):
        """Newton-Raphson from lengths[slot] to lengths[slot], nothing read back."""

        class Newton(C.Structure):
            _fields_ = [("min_len", C.c_double), ("max_len", C.c_double), ("xacc", C.c_double),
                        ("guess", C.c_double), ("max_steps", C.c_int32), ("reserved", C.c_int32)]
        nr = Newton(min_len, max_len, min_len if xacc is None else xacc, 0.0, max_steps, 0)
        f = self._f("optimize_branch_slot")
        f.argtypes = [C.c_void_p, End, End, C.c_void_p, C.c_int32]
        self.last_status = self._check(f(self.ctx, End(*dad), End(*node), C.byref(nr), C.c_int32(slot)))

    def evaluate(self, entries, dad, node, length, want_lnl=True, want_derv=True):
        nodes, nn, kids, nk = pack_traversal(entries)
        lnl, df, ddf = C.c_double(), C.c_double(), C.c_double()
        f = self._f("evaluate")
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, End, End, C.c_double,
                      C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
        self.last_status = self._check(
            f(self.ctx, nodes, nn, kids, nk, End(*dad), End(*node), length,
              C.byref(lnl) if want_lnl else None, C.byref(df) if want_derv else None,
              C.byref(ddf) if want_derv else None))
        return lnl.value, df.value, ddf.value

    def evaluate_async(self, packed, dad, node, length, want_lnl=True, want_derv=True):
        """Enqueue a whole evaluation; ``packed`` = pack_traversal(entries) (reusable).  Results stay
        on the device (result_tensor) until read_results()."""
        nodes, nn, kids, nk = packed
        f = self._f("evaluate_async")
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, End, End, C.c_double,
                      C.c_int, C.c_int]
        return self._check(f(self.ctx, nodes, nn, kids, nk, End(*dad), End(*node), length,
                             int(want_lnl), int(want_derv)))

    def read_results(self):
        lnl, df, ddf = C.c_double(), C.c_double(), C.c_double()
        self.last_status = self._check(self._f("read_results")(self.ctx, C.byref(lnl), C.byref(df), C.byref(ddf)))
        return lnl.value, df.value, ddf.value

    def set_stream(self, cuda_stream_ptr):
        """Run on a caller-owned stream (an int cudaStream_t, e.g. torch.cuda.current_stream().cuda_stream)."""
        return self._check(self._f("set_stream")(self.ctx, C.c_void_p(cuda_stream_ptr)))

    def peer_handle(self):
        """64-byte CUDA IPC handle of this rank's exchange buffer (b200lk_peer_handle)."""
        buf = C.create_string_buffer(64)
        self._check(self._f("peer_handle")(self.ctx, buf))
        return buf.raw

    def peer_connect(self, rank, world, handles):
        """handles: bytes, world x 64, rank-ordered.  From then on evaluate_async / evaluate leave the sums over ALL ranks in
        every rank's result buffer: the finishing kernel exchanges them over NVLink peer memory (b200lk_peer_connect)."""
        assert len(handles) == 64 * world
        self._check(self._f("peer_connect")(self.ctx, C.c_int32(rank), C.c_int32(world), C.c_char_p(handles)))

    def peer_disconnect(self):
        self._check(self._f("peer_disconnect")(self.ctx))

    def connect_peers(self, dist):
        """Gather every rank's handle over torch.distributed and connect (one process per GPU)."""
        import torch
        rank, world = dist.get_rank(), dist.get_world_size()
        dev = "cuda:%d" % self.dims.device if dist.get_backend() == "nccl" else "cpu"
        mine = torch.frombuffer(bytearray(self.peer_handle()), dtype=torch.uint8).to(dev)
        allh = [torch.empty(64, dtype=torch.uint8, device=dev) for _ in range(world)]
        dist.all_gather(allh, mine)
        self.peer_connect(rank, world, b"".join(bytes(t.cpu().numpy().tobytes()) for t in allh))

    def set_profiling(self, on=True):
        return self._check(self._f("set_profiling")(self.ctx, int(on)))

    def get_profile(self, reset=True):
        ms, n = C.c_double(), C.c_int64()
        self._check(self._f("get_profile")(self.ctx, C.byref(ms), C.byref(n), int(reset)))
        return ms.value, n.value

    def result_tensor(self):
        """The device result buffer [lnL, df, ddf] as a torch tensor aliasing device memory, so that
        pattern shards can be all-reduced in place (NCCL) on the context's stream."""
        import torch
        p = C.c_void_p()
        self._check(self._f("result_device_ptr")(self.ctx, C.byref(p)))

        class _Buf:
            __cuda_array_interface__ = {"shape": (3,), "typestr": "<f8", "data": (p.value, False), "version": 3}
        return torch.as_tensor(_Buf(), device="cuda:%d" % self.dims.device)

    # -- readers -------------------------------------------------------------
    def fetch_pattern_lh(self):
        out = np.empty(self.dims.nptn)
        self._check(self._f("fetch_pattern_lh")(self.ctx, _dp(out)))
        return out

    def fetch_pattern_lh_cat(self):
        out = np.empty((self.dims.nptn, self.ncat_mix))
        self._check(self._f("fetch_pattern_lh_cat")(self.ctx, _dp(out)))
        return out

    def fetch_scale_num(self, slot):
        shape = (self.dims.nptn, self.ncat_mix) if self.dims.safe_numeric else (self.dims.nptn,)
        out = np.empty(shape, dtype=np.uint8)
        self._check(self._f("fetch_scale_num")(self.ctx, C.c_int32(slot), _u8p(out)))
        return out

    def fetch_partial(self, slot):
        out = np.empty((self.dims.nptn, self.ncat_mix, self.dims.nstates))
        self._check(self._f("fetch_partial")(self.ctx, C.c_int32(slot), _dp(out)))
        return out

    def fetch_theta(self):
        th = np.empty((self.dims.nptn, self.ncat_mix, self.dims.nstates))
        bs = np.empty(self.dims.nptn)
        self._check(self._f("fetch_theta")(self.ctx, _dp(th), _dp(bs)))
        return th, bs

    def upload_boot_samples(self, samples):
        """UFBoot: [nsamples][nptn] float32 resampled pattern frequencies, resident on the device."""
        a = np.ascontiguousarray(samples, dtype=np.float32)
        assert a.ndim == 2 and a.shape[1] == self.dims.nptn
        self._check(self._f("upload_boot_samples")(self.ctx, a.ctypes.data_as(C.c_void_p), C.c_int32(a.shape[0])))
        self._nboot = a.shape[0]

    def rell(self):
        """RELL log-likelihood of every resample from the device-resident per-pattern lnL of the last evaluation."""
        out = np.empty(self._nboot)
        self._check(self._f("rell")(self.ctx, _dp(out)))
        return out

    def shard_info(self):
        """-> [(device, begin, end), ...]: the pattern range every shard holds (one entry for an ordinary context)."""
        n = C.c_int32()
        self._check(self._f("shard_info")(self.ctx, C.byref(n), None, None, None))
        dev, b, e = (C.c_int32 * n.value)(), (C.c_int64 * n.value)(), (C.c_int64 * n.value)()
        self._check(self._f("shard_info")(self.ctx, C.byref(n), dev, b, e))
        return [(dev[i], b[i], e[i]) for i in range(n.value)]

    def counters(self):
        a, b = C.c_int64(), C.c_int64()
        self._check(self._f("get_counters")(self.ctx, C.byref(a), C.byref(b)))
        return a.value, b.value

    def min_traffic(self):
        """Minimum DRAM bytes of the partial updates launched since the last reset (b200lk_get_min_traffic)."""
        b = C.c_int64()
        self._check(self._f("get_min_traffic")(self.ctx, C.byref(b)))
        return b.value

    def reset_counters(self):
        self._check(self._f("reset_counters")(self.ctx))

    def synchronize(self):
        if self.prefix == "b200lk_":
            self._check(self.lib.b200lk_synchronize(self.ctx))


class Batch:
    """Engines of one supermatrix that share dimensions and topology, evaluated with batched launches
    (include/b200lk.h, b200lk_batch_*).  Raises B200LKError(status -5) for configurations a batch cannot express."""

    def __init__(self, engines):
        self.engines = list(engines)
        self.n = len(self.engines)
        self.lib = self.engines[0].lib
        arr = (C.c_void_p * self.n)(*[e.ctx for e in self.engines])
        self.h = C.c_void_p()
        f = self.lib.b200lk_batch_create
        f.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_int32]
        rc = f(C.byref(self.h), arr, self.n)
        if rc < 0:
            self._raise(rc)

    def _raise(self, rc):
        msg = ""
        if self.h:
            self.lib.b200lk_batch_last_error.restype = C.c_char_p
            self.lib.b200lk_batch_last_error.argtypes = [C.c_void_p]
            msg = (self.lib.b200lk_batch_last_error(self.h) or b"").decode()
        raise B200LKError("b200lk_batch call failed with status %d %s" % (rc, msg))

    def close(self):
        if self.h:
            self.lib.b200lk_batch_destroy.argtypes = [C.c_void_p]
            self.lib.b200lk_batch_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def evaluate(self, entries, child_len, dad, node, branch_len, want_derv=True):
        """entries: the shared traversal (lengths inside are ignored); child_len [n][nchildren]; branch_len [n].
        -> (lnl[n], df[n], ddf[n])"""
        nodes, nn, kids, nc = pack_traversal(entries)
        child_len = np.ascontiguousarray(child_len, dtype=np.float64).reshape(self.n, max(nc, 0))
        branch_len = np.ascontiguousarray(branch_len, dtype=np.float64).reshape(self.n)
        lnl, df, ddf = (np.zeros(self.n) for _ in range(3))
        f = self.lib.b200lk_batch_evaluate
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.POINTER(C.c_double), End, End,
                      C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                      C.POINTER(C.c_double)]
        rc = f(self.h, nodes, nn, kids, nc, _dp(child_len), End(*dad), End(*node), _dp(branch_len), int(want_derv),
               _dp(lnl), _dp(df), _dp(ddf))
        if rc < 0:
            self._raise(rc)
        return lnl, df, ddf

    def build_theta(self, dad, node):
        f = self.lib.b200lk_batch_build_theta
        f.argtypes = [C.c_void_p, End, End]
        rc = f(self.h, End(*dad), End(*node))
        if rc < 0:
            self._raise(rc)

    def derv(self, branch_len):
        branch_len = np.ascontiguousarray(branch_len, dtype=np.float64).reshape(self.n)
        df, ddf = np.zeros(self.n), np.zeros(self.n)
        f = self.lib.b200lk_batch_derv
        f.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
        rc = f(self.h, _dp(branch_len), _dp(df), _dp(ddf))
        if rc < 0:
            self._raise(rc)
        return df, ddf

    def lnl_from_buffer(self, branch_len):
        branch_len = np.ascontiguousarray(branch_len, dtype=np.float64).reshape(self.n)
        lnl = np.zeros(self.n)
        f = self.lib.b200lk_batch_lnl_from_buffer
        f.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        rc = f(self.h, _dp(branch_len), _dp(lnl))
        if rc < 0:
            self._raise(rc)
        return lnl


def engine_for(aln, model, rates, props, *, nslots, p_invar=0.0, safe_numeric=False, device=0,
               lib_path=LIB_PATH, prefix="b200lk_", unobserved=None, fused=False, asc_holder=False, devices=None):
    """Create a context for a workload.Alignment + workload.EigenModel and load
    tips, pattern weights and the model (the one-time uploads of
    initializeAllPartialLh / computeTipPartialLikelihood)."""
    from . import workload as W
    eng = Engine(ntaxa=aln.ntaxa, nstates=aln.nstates, ncat=len(rates), nmix=model.nmix, nptn=aln.nptn,
                 nslots=nslots, n_state_codes=aln.state_unknown + 1, safe_numeric=safe_numeric,
                 device=device, lib_path=lib_path, prefix=prefix, fused_mix_rate=fused,
                 n_unobserved=0 if unobserved is None else np.asarray(unobserved).shape[1], devices=devices)
    eng.upload_tips(aln.states, aln.ptn_freq, unobserved)
    if asc_holder:
        eng.set_asc(1)
    inv = W.ptn_invar(aln, model, p_invar)
    if unobserved is not None:
        # +ASC: the unobserved patterns follow the observed ones; their ptn_invar entries are the reference's
        # "dummy values" p_invar (tree/phylotreesse.cpp:607-609)
        inv = np.concatenate([inv, np.full(np.asarray(unobserved).shape[1], p_invar)])
    eng.set_ptn_invar(inv)
    scale = 1.0 - p_invar
    eng.set_model(model.eval, model.evec, model.inv_evec, np.asarray(rates) , np.asarray(props) * scale,
                  model.mixw, W.tip_partial_lh(model))
    return eng
