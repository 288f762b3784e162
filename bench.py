#!/usr/bin/env python
"""bench.py -- the tree-likelihood hot path on B200, one JSON line per run.

A *step* is one full evaluation of the workload through the C ABI (include/b200lk.h):
every partial-likelihood vector stale -> post-order update of all T-2 internal nodes ->
lnL at the root branch -> theta -> (df, ddf): SURVEY.md section 8(d), metric (1)+(2).

  value     pattern-node likelihood updates/s, device-timed (CUDA events on the launching stream), inputs resident
            in HBM, results left on the device (b200lk_evaluate_async)
  e2e       the same through the blocking call a host caller makes per model probe: b200lk_set_model (eigen-system +
            rates H2D) + b200lk_evaluate (traversal descriptors and branch lengths H2D, lnL/df/ddf D2H)
  roofline  the partial-update kernel: MINIMUM DRAM bytes of the node programs actually launched
            (b200lk_get_min_traffic) / its CUDA-event time inside the timed region, against MEASURED_PEAKS.json.
            frac <= 1 by construction.  SURVEY.md 8(d)'s level-by-level byte count is reported beside it as
            `algorithmic_bytes_per_step` and `fusion_dividend` (how many of those bytes the kernel never moves).
  parity    GPU results against the reference's own AVX+FMA kernel (oracle/_ref) on the SAME patterns, in this run;
            the run exits non-zero when it exceeds 1e-9
  cpu_baseline  the reference kernel timed on the host cores on the full workload (same_config true)
  shapes    sub-results for the other BASELINE configs (cfg3, cfg4 at full size with -mem, cfg5, cfg2 branch sweep)

N > 1 (torchrun): the SAME alignment on every rank, sharded by site pattern (b200::shardPatterns): strong scaling of the
named shape; the only exchange is an NCCL all-reduce of [lnL, df, ddf] in place on the device result buffer, and the
reduced values must equal the reference's.  The weak-scaling figure (the whole alignment on every GPU) is a secondary key.

`--impl reference` times the CPU implementation only (oracle/_ref when built, else the oracle port).
"""
import ctypes
import os

# The reference's OpenMP loops run with their threads pinned to cores (OMP_PROC_BIND=close, OMP_PLACES=cores).  libgomp
# reads those variables once, when it is loaded, and at that moment also binds the CALLING thread to the first place --
# which would put the main thread of every torchrun rank, and the NCCL threads that inherit its mask, on core 0.  So: load
# the OpenMP runtime here, under the binding variables, and give the main thread its full affinity mask back at once.
# OpenMP worker threads are pinned individually when a team starts; nothing else in the process is.
_AFFINITY = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else None
os.environ.setdefault("OMP_PROC_BIND", "close")
os.environ.setdefault("OMP_PLACES", "cores")
try:
    ctypes.CDLL("libgomp.so.1", mode=ctypes.RTLD_GLOBAL)
except OSError:
    pass
if _AFFINITY is not None:
    os.sched_setaffinity(0, _AFFINITY)

import argparse  # noqa: E402
import json  # noqa: E402
import sys  # noqa: E402
import threading  # noqa: E402
import time  # noqa: E402

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "pattern_node_likelihood_updates_per_sec"
UNIT = "updates/s"
TOL = 1e-9
T_START = time.perf_counter()

WORKLOADS = {
    # name: (ntaxa, npatterns, nstates, ncat, nmix, alpha, seed, description)
    "cfg2": (200, 1_000_000, 4, 4, 1, 0.5, 1, "synthetic DNA 200 taxa x 1M patterns, GTR+G4, full traversal + lnL + derv"),
    "cfg3": (500, 200_000, 20, 4, 1, 0.7, 2, "synthetic protein 500 taxa x 200k patterns, LG+G4"),
    # cfg4 is 100k patterns: 376 GB of partial vectors with one slot per node.  One GPU holds it with -mem
    # (--mem-save 0); without it the shape is 100k patterns over 8 GPUs, 12.5k per GPU
    "cfg4": (100, 12_500, 20, 4, 60, 0.7, 3, "synthetic protein 100 taxa x 100k patterns / 8 GPUs (12.5k per GPU), LG+C60+G4"),
    "cfg4full": (100, 100_000, 20, 4, 60, 0.7, 3, "synthetic protein 100 taxa x 100k patterns, LG+C60+G4"),
    "cfg5": (100, 1000, 4, 4, 1, 0.5, 4, "partitioned supermatrix (see measure_partitioned)"),
    "small": (50, 100_000, 4, 4, 1, 0.5, 5, "synthetic DNA 50 taxa x 100k patterns, GTR+G4 (smoke size)"),
    # not a BASELINE config: the 61-state codon case north_star names for the tensor path
    "codon": (100, 50_000, 61, 4, 1, 0.5, 6, "synthetic codon 100 taxa x 50k patterns, 61-state reversible +G4"),
}


def model_for(n, nmix, seed):
    """The substitution model BASELINE.json names for the workload: GTR for DNA, LG for 20-state single-class,
    LG+C60 (60 profile classes over LG exchangeabilities) for the 60-class mixture; anything else a seeded random
    reversible model."""
    from iqtree2_b200 import workload as W
    if n == 4 and nmix == 1:
        return W.gtr_model(), "GTR{1,2,1.5,1.2,2.5,1}+F{0.3,0.2,0.2,0.3}"
    if n == 20 and nmix == 1 and hasattr(W, "lg_model"):
        return W.lg_model(), "LG"
    if n == 20 and nmix == 60 and hasattr(W, "lg_c60_model"):
        return W.lg_c60_model(), "LG+C60"
    return W.random_reversible_model(n, seed + 100, nmix), "random reversible %d-state model, %d classes" % (n, nmix)


_WL_CACHE = {}


def build_workload(name, npatterns=None):
    """(tree, model, rates, props, alignment, model name); the last workload built is kept (the cfg2 sub-results reuse
    the alignment of the headline run instead of simulating it again)."""
    from iqtree2_b200 import workload as W
    T, N, n, ncat, nmix, alpha, seed, _ = WORKLOADS[name]
    if npatterns:
        N = npatterns
    if (name, N) in _WL_CACHE:
        return _WL_CACHE[(name, N)]
    _WL_CACHE.clear()
    tree = W.random_tree(T, seed)
    model, model_name = model_for(n, nmix, seed)
    rates, props = W.discrete_gamma(alpha, ncat)
    aln = W.simulate_alignment(tree, model, rates, props, N, seed + 1000, compress=False)
    _WL_CACHE[(name, N)] = (tree, model, rates, props, aln, model_name)
    return _WL_CACHE[(name, N)]


def traversal_flops(tr, N, nstates, ncat_mix):
    """Algorithmic flops of the partial updates of one traversal:
    F_K = C * (2 n^2 (k_int + 1) + n) per pattern for a bifurcating node with k_int internal children."""
    total = 0
    for e in tr:
        k = sum(1 for ch in e[3] if ch[0] == 1)
        total += ncat_mix * (2 * nstates * nstates * (k + 1) + nstates)
    return total * N


def traversal_bytes(tr, N, nstates, ncat_mix, safe=False):
    """Algorithmic bytes of the partial updates of one traversal:
    B_K = 8*block*(1+k_int) + S*(1+k_int) + s*(2-k_int) per pattern for a bifurcating node with
    k_int internal children."""
    block = nstates * ncat_mix
    S = ncat_mix if safe else 1
    total = 0
    kinds = {0: 0, 1: 0, 2: 0}
    for e in tr:
        k = sum(1 for ch in e[3] if ch[0] == 1)
        nleaf = len(e[3]) - k
        total += 8 * block * (1 + k) + S * (1 + k) + nleaf
        kinds[min(k, 2)] += 1
    return total * N, kinds


def peaks():
    """(hbm GB/s, source), (fp64 DMMA TFLOP/s, source).  HBM: MEASURED_PEAKS.json (driver-written) else the profiling
    recipe's fallback.  FP64 tensor: MEASURED_PEAKS.json has no fp64 entry; the DMMA m8n8k4 peak measured by
    tools/fp64_peak.cu on this pool's B200 (profiles/r1g_fp64_peak.json) is used and named."""
    hbm, hbm_src = 6650.0, "fallback 6650 GB/s (B200_PROFILING.md; MEASURED_PEAKS.json absent)"
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        try:
            hbm, hbm_src = float(json.load(open(pk))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
        except Exception:
            pass
    f64, f64_src = 37.1, "fallback 37.1 TFLOP/s: DMMA m8n8k4 peak measured once with tools/fp64_peak.cu"
    fp = os.path.join(ROOT, "profiles", "r1g_fp64_peak.json")
    if os.path.exists(fp):
        try:
            f64 = float(json.load(open(fp))["dmma_tflops"])
            f64_src = ("profiles/r1g_fp64_peak.json (DMMA m8n8k4 peak measured with tools/fp64_peak.cu; "
                       "MEASURED_PEAKS.json has no fp64 entry)")
        except Exception:
            pass
    return (hbm, hbm_src), (f64, f64_src)


_HOST_CORES = None


def host_cores():
    """(threads used by the CPU legs, physical cores, logical CPUs).  The reference gets every logical CPU the
    process may run on (`-T N` with N = all of them), pinned (OMP_PROC_BIND=close, OMP_PLACES=cores).  Taken once,
    before an OpenMP runtime binds the main thread to its own place."""
    global _HOST_CORES
    if _HOST_CORES is None:
        _HOST_CORES = _host_cores()
    return _HOST_CORES


def _host_cores():
    logical = len(_AFFINITY) if _AFFINITY is not None else (os.cpu_count() or 1)
    physical = None
    try:
        import psutil
        physical = psutil.cpu_count(logical=False)
    except Exception:
        pass
    return logical, physical, os.cpu_count() or logical


def host_ram_available():
    try:
        import psutil
        return int(psutil.virtual_memory().available)
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.max_mhz = index, False, [], set(), None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                     nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
            while not self.stop_flag:
                self.sm.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
                time.sleep(0.02)
        except Exception as ex:  # no NVML: report that instead of guessing
            self.reasons.add("nvml_unavailable:%s" % type(ex).__name__)

    def finish(self):
        self.stop_flag = True
        self.join()
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    if a.size == 0:
        return 0.0
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.maximum(np.abs(a), np.abs(b)))))


# ---- the CPU leg: the reference's own kernel (oracle/_ref) on the host cores -----------------------------------------
class CpuLeg:
    """The reference's AVX+FMA kernel with OpenMP (oracle/_ref, kind "reference") on patterns [0, nptn) of the workload,
    or the C restatement (kind "port") where that library was not built.  One evaluation = all partials stale -> lnL
    at the branch of taxon 0 -> theta -> df/ddf: exactly the GPU step.  Used for `parity`, `cpu_baseline` and
    `--impl reference` only."""

    def __init__(self, tree, model, rates, props, aln, nptn=None, threads=None):
        import oracle
        from iqtree2_b200 import workload as W
        from iqtree2_b200.capi import INTERNAL, LEAF
        self.threads = threads or host_cores()[0]
        block = aln.nstates * len(rates) * model.nmix
        want = aln.nptn if not nptn else min(nptn, aln.nptn)
        # the reference keeps one partial vector per internal node on the host: shrink the prefix if RAM is short
        need = lambda k: (tree.ntaxa - 2) * k * block * 8 * 1.08 + k * tree.ntaxa * 5 + (1 << 28)
        avail = host_ram_available()
        self.ram_limited = False
        while avail is not None and want > 1000 and need(want) > 0.85 * avail:
            want = max(1000, want // 2)
            self.ram_limited = True
        self.nptn, self.full = want, want == aln.nptn
        self.sub = aln if self.full else W.Alignment(aln.nstates, aln.names, np.ascontiguousarray(aln.states[:, :want]),
                                                     np.ascontiguousarray(aln.ptn_freq[:want]))
        self.tree = tree
        c = tree.adj[0][0]
        self.tr = W.traversal(tree, c, 0)
        if oracle.ref_available():
            self.kind = "reference"
            self.what = "reference AVX+FMA kernel (-lk FMA), OpenMP"
            self.ref = oracle.RefRun(tree, self.sub, model, rates, props, kernel=oracle.LK_AVX_FMA, threads=self.threads)
        else:
            self.kind = "port"
            self.what = "C restatement of the reference kernel, OpenMP"
            self.eng = oracle.oracle_engine(self.sub, model, rates, props, nslots=tree.ntaxa)
            self.eng._f("set_threads")(self.eng.ctx, self.threads)
            self.args = (self.tr, (LEAF, 0), (INTERNAL, self.tr[-1][2]), tree.get_length(0, c))

    def evaluate(self):
        if self.kind == "reference":
            lnl = self.ref.lnl_at_leaf(0)
            df, ddf = self.ref.derv()
            return lnl, df, ddf
        return self.eng.evaluate(*self.args)

    def pattern_lh(self):
        return self.ref.pattern_lh() if self.kind == "reference" else self.eng.fetch_pattern_lh()

    def top_scale(self):
        """scale counters of the traversal's last node (the whole tree but taxon 0)"""
        if self.kind == "reference":
            return self.ref.fetch_branch(range(1, self.tree.ntaxa), 0)[1]
        return self.eng.fetch_scale_num(self.tr[-1][2])

    def timed(self, steps, warmup):
        for _ in range(warmup):
            res = self.evaluate()
        t0 = time.perf_counter()
        for _ in range(steps):
            res = self.evaluate()
        dt = (time.perf_counter() - t0) / max(1, steps)
        return dt, res

    def baseline(self, dt, steps, total_patterns):
        logical, physical, _ = host_cores()
        return {"value": len(self.tr) * self.nptn / dt, "unit": UNIT, "cores": self.threads, "kind": self.kind,
                "physical_cores": physical, "logical_cpus": logical,
                "pinning": "OMP_PROC_BIND=%s OMP_PLACES=%s" % (os.environ.get("OMP_PROC_BIND"), os.environ.get("OMP_PLACES")),
                "same_config": self.full,
                "sample": "%s, %d threads; %s%d of %d patterns, same tree/model, %d full evaluations"
                          % (self.what, self.threads, "all " if self.full else "first ", self.nptn, total_patterns, steps),
                "ms_per_eval": dt * 1e3, "evals_per_sec": 1.0 / dt}

    def close(self):
        if self.kind == "reference":
            self.ref.close()
        else:
            self.eng.close()


def gpu_eval_patterns(tree, model, rates, props, aln, nptn, device, mem_safe=False):
    """(lnl, df, ddf), per-pattern lnL and top-node scale counters of the CUDA path on patterns [0, nptn)."""
    from iqtree2_b200 import workload as W
    from iqtree2_b200.capi import INTERNAL, LEAF, engine_for
    sub = aln if nptn == aln.nptn else W.Alignment(aln.nstates, aln.names, np.ascontiguousarray(aln.states[:, :nptn]),
                                                  np.ascontiguousarray(aln.ptn_freq[:nptn]))
    eng = engine_for(sub, model, rates, props, nslots=tree.ntaxa, device=device)
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    res = eng.evaluate(tr, (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c))
    out = res, eng.fetch_pattern_lh(), eng.fetch_scale_num(tr[-1][2])
    eng.close()
    return out


def parity_block(gpu_res, cpu_res, gpu_plh, cpu_plh, gpu_scale, cpu_scale, kind, extra=None):
    """GPU against the CPU leg on the same patterns: north_star's bar (scale counts bit-exact; lnL, df, ddf and
    per-pattern lnL within 1e-9 relative).  The per-pattern arrays cover the patterns both sides hold."""
    k = min(len(gpu_plh), len(cpu_plh))
    p = {"against": kind, "tol": TOL,
         "lnl_rel": rel(gpu_res[0], cpu_res[0]), "df_rel": rel(gpu_res[1], cpu_res[1]), "ddf_rel": rel(gpu_res[2], cpu_res[2]),
         "pattern_lh_rel": rel(gpu_plh[:k], cpu_plh[:k]), "patterns_compared_one_by_one": int(k),
         "scale_counts_equal": bool(np.array_equal(np.asarray(gpu_scale)[:k], np.asarray(cpu_scale)[:k])),
         "scale_count_max": int(np.max(cpu_scale)) if len(cpu_scale) else 0,
         "gpu": [float(x) for x in gpu_res], "cpu": [float(x) for x in cpu_res]}
    if extra:
        p.update(extra)
    p["ok"] = bool(p["scale_counts_equal"] and all(v <= TOL for kk, v in p.items() if kk.endswith("_rel")))
    return p


_STDOUT_FD = None


def own_stdout():
    """stdout carries exactly ONE JSON line: libraries that write banners to fd 1 (NCCL's version line, compiler
    chatter of an on-demand build) are sent to stderr for the life of the process; emit() writes to the real stdout."""
    global _STDOUT_FD
    if _STDOUT_FD is None:
        sys.stdout.flush()
        _STDOUT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _STDOUT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_STDOUT_FD, data)


def log(msg):
    sys.stderr.write("[bench %6.1fs] %s\n" % (time.perf_counter() - T_START, msg))
    sys.stderr.flush()


class Dist:
    """torchrun plumbing: one process per GPU over NCCL (RANK / LOCAL_RANK / WORLD_SIZE from the environment)."""

    def __init__(self):
        import torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback on this path")
        torch.cuda.set_device(self.local)
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # NCCL's version banner goes to stdout: keep stdout to the one JSON line
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist

    def sync(self):
        import torch
        if self.dist:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max(self, *vals):
        import torch
        t = torch.tensor(vals, dtype=torch.float64, device="cuda")
        if self.dist:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t.cpu()]

    def close(self):
        if self.dist:
            self.dist.destroy_process_group()


def timed_engine_run(D, eng, packed, dad, node, L, model_args, steps, warmup, collective="peer"):
    """W warm-up + `steps` device-timed evaluations of `eng` (CUDA events on the launching stream, barrier +
    synchronize on both sides), then the same through the blocking host-facing calls (e2e).  Returns a dict."""
    import torch
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    eng.set_stream(stream.cuda_stream)  # one stream for the engine, NCCL and the timing events
    res_t = eng.result_tensor()
    # the sums over the ranks: exchanged inside the finishing kernel over NVLink peer memory (b200lk_peer_connect), or
    # with --collective nccl by an ncclAllReduce of the 3 doubles behind it
    use_nccl = D.dist is not None and collective == "nccl"
    peer_note = None
    if D.dist is not None and not use_nccl:
        try:
            eng.connect_peers(D.dist)
            ok = 1
        except Exception as ex:  # no CUDA IPC / peer access on this box: every rank falls back to NCCL together
            peer_note, ok = "%s: %s" % (type(ex).__name__, ex), 0
        import torch
        flag = torch.tensor([ok], device="cuda")
        D.dist.all_reduce(flag, op=D.dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            if ok:
                eng.peer_disconnect()
            use_nccl = True
            log("peer exchange unavailable (%s): using ncclAllReduce" % peer_note)

    def step_device():
        eng.evaluate_async(packed, dad, node, L)
        if use_nccl:
            D.dist.all_reduce(res_t)

    def step_e2e():
        eng.set_model(*model_args)
        eng.evaluate_async(packed, dad, node, L)
        if use_nccl:
            D.dist.all_reduce(res_t)
        return eng.read_results()

    for _ in range(max(3, warmup)):
        step_device()
    D.sync()
    eng.synchronize()
    eng.get_profile(reset=True)
    eng.reset_counters()
    sampler = ClockSampler(D.local)
    sampler.start()
    eng.set_profiling(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    D.sync()
    e0.record(stream)
    for _ in range(steps):
        step_device()
    e1.record(stream)
    D.sync()
    ms = e0.elapsed_time(e1)
    eng.set_profiling(False)
    partial_ms, partial_launches = eng.get_profile(reset=True)
    _, launches = eng.counters()
    min_bytes = eng.min_traffic()
    local_res = None
    if use_nccl:  # this rank's own sums, for the check that the all-reduce really summed finished kernels
        eng.evaluate_async(packed, dad, node, L)
        local_res = eng.read_results()
    # ---- end to end: host buffers in, scalars out, every step ----
    for _ in range(2):
        step_e2e()
    D.sync()
    t0 = time.perf_counter()
    for _ in range(steps):
        result = step_e2e()
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.finish()
    coll_us = None
    if D.dist:  # the collective alone: all-reduce of the 3 result doubles, back to back
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        scratch = res_t.clone()
        for _ in range(20):
            D.dist.all_reduce(scratch)
        torch.cuda.synchronize()
        c0.record(stream)
        for _ in range(200):
            D.dist.all_reduce(scratch)
        c1.record(stream)
        torch.cuda.synchronize()
        coll_us = c0.elapsed_time(c1) * 1e3 / 200
        # leave the true (summed over ranks) result of one more evaluation in the buffer
        result = step_e2e()
    ms, e2e_ms = D.max(ms, e2e_ms)
    torch.cuda.set_stream(torch.cuda.default_stream())
    eng.set_stream(None)
    return {"ms": ms, "e2e_ms": e2e_ms, "partial_ms": partial_ms, "partial_launches": partial_launches,
            "launches": launches, "min_bytes": min_bytes, "result": result, "local_result": local_res,
            "clocks": clocks, "collective_us": coll_us, "used_nccl": use_nccl, "peer_note": peer_note}


def timed_sharded_run(eng, packed, dad, node, L, model_args, steps, warmup, ndev):
    """The same two regions for a b200lk_create_sharded context: ONE process drives every device.  A step enqueues the
    evaluation on all shards and then reads the per-shard sums back (added on the host in device-index order), so
    every step ends with a blocking read and the wall clock around K steps is the device time of the slowest shard
    plus the fan-out; no stream or event spans the devices."""
    import torch

    def sync_all():
        eng.synchronize()
        for d in range(ndev):
            torch.cuda.synchronize(d)

    def step():
        eng.evaluate_async(packed, dad, node, L)
        return eng.read_results()

    def step_e2e():
        eng.set_model(*model_args)
        eng.evaluate_async(packed, dad, node, L)
        return eng.read_results()

    for _ in range(max(3, warmup)):
        step()
    sync_all()
    eng.get_profile(reset=True)
    eng.reset_counters()
    samplers = [ClockSampler(d) for d in range(ndev)]
    for sp in samplers:
        sp.start()
    eng.set_profiling(True)
    sync_all()
    t0 = time.perf_counter()
    for _ in range(steps):
        result = step()
    sync_all()
    ms = (time.perf_counter() - t0) * 1e3
    eng.set_profiling(False)
    partial_ms, partial_launches = eng.get_profile(reset=True)
    _, launches = eng.counters()
    min_bytes = eng.min_traffic()
    for _ in range(2):
        step_e2e()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(steps):
        result = step_e2e()
    sync_all()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    clocks = [sp.finish() for sp in samplers]
    worst = min(clocks, key=lambda c: c["sm_mhz"] or 1e9)
    worst = dict(worst, reasons=sorted(set(r for c in clocks for r in c["reasons"])), per_device_sm_mhz=[c["sm_mhz"] for c in clocks])
    return {"ms": ms, "e2e_ms": e2e_ms, "partial_ms": partial_ms, "partial_launches": partial_launches,
            "launches": launches, "min_bytes": min_bytes, "result": result, "local_result": None,
            "clocks": worst, "collective_us": None}


def attach_ncu_traffic(roof, name, full_size):
    """roofline.traffic: dram__bytes_read + dram__bytes_write of the dominant kernel PER LAUNCH from the committed ncu pass
    of this workload (profiles/traffic.json, tagged with the commit it was captured at) -- only when this run is the
    size that pass measured; otherwise null."""
    roof.pop("_workload_key", None)
    if isinstance(roof.get("hbm"), dict):
        roof["hbm"].pop("_workload_key", None)
    tr_file = os.path.join(ROOT, "profiles", "traffic.json")
    if not full_size or not os.path.exists(tr_file):
        return
    try:
        t = json.load(open(tr_file)).get(name)
    except Exception:
        return
    if t:
        roof["traffic"] = t["dram_bytes_per_launch"]
        roof["traffic_source"] = {k: t[k] for k in ("dram_bytes_per_step", "launches_per_step", "commit", "source") if k in t}


def roofline_of(n, ncm, tr, N_local, r, steps):
    """roofline object of the partial-update kernel(s) from one timed_engine_run on N_local patterns."""
    (hbm, hbm_src), (f64, f64_src) = peaks()
    alg_bytes, kinds = traversal_bytes(tr, N_local, n, ncm)
    alg_flops = traversal_flops(tr, N_local, n, ncm)
    kms = r["partial_ms"] / max(1, steps)
    min_b = r["min_bytes"] / max(1, steps)
    gbs = min_b / (kms * 1e-3) / 1e9 if kms > 0 else None
    tfl = alg_flops / (kms * 1e-3) / 1e12 if kms > 0 else None
    hbm_obj = {"bound": "hbm", "achieved": gbs, "peak": hbm, "peak_source": hbm_src, "unit": "GB/s",
               "frac": gbs / hbm if gbs else None, "traffic": None,
               "min_dram_bytes_per_step": int(min_b),
               "bytes_definition": "b200lk_get_min_traffic: from the node programs launched -- every vector written once, "
                                   "every child a launch must fetch from HBM, one byte per tip state",
               "algorithmic_bytes_per_step": int(alg_bytes),
               "fusion_dividend": alg_bytes / min_b if min_b else None,
               "kernel_ms_per_step": kms, "launches_per_step": r["partial_launches"] / max(1, steps)}
    if n == 4:
        hbm_obj["kernel"] = "s4_traverse_kernel (4-state partial update, whole traversal)"
    else:
        hbm_obj["kernel"] = "partial update (pruning step), level-batched"
    hbm_obj["_workload_key"] = None
    if n < 9:
        return hbm_obj, kinds
    # dense paths: the FP64 tensor pipe (DMMA) bounds the kernel next to HBM (20 states: AI ~ 5 flop/B, at the ridge)
    ten = {"bound": "tensor",
           "kernel": "dense20_rp_kernel + dense20_tt_kernel (20-state partial update, FP64 DMMA, level-batched)" if n == 20
           else "denseN_kernel (%d-state partial update, FP64 DMMA on 8-state tiles, level-batched)" % n,
           "achieved": tfl, "peak": f64, "peak_source": f64_src, "unit": "TFLOP/s", "frac": tfl / f64 if tfl else None,
           "traffic": None, "algorithmic_flops_per_step": int(alg_flops), "kernel_ms_per_step": kms,
           "launches_per_step": r["partial_launches"] / max(1, steps),
           "hbm": {k: hbm_obj[k] for k in ("achieved", "peak", "unit", "frac", "min_dram_bytes_per_step")}}
    return ten, kinds


def measure_alignment(args, D, name, steps, warmup, want_cpu_baseline, cpu_patterns=0, weak_too=True):
    """One alignment, sharded by site pattern over the ranks (strong scaling of the named shape).  Returns the JSON
    line as a dict on rank 0 (None elsewhere)."""
    import torch
    from iqtree2_b200 import multigpu as MG, workload as W
    from iqtree2_b200.capi import INTERNAL, LEAF, engine_for, pack_traversal
    T, N0, n, ncat, nmix, *_ = WORKLOADS[name]
    t0 = time.perf_counter()
    tree, model, rates, props, aln, model_name = build_workload(name, args.patterns or None)
    N = aln.nptn
    gen_s = time.perf_counter() - t0
    log("%s: alignment %d x %d simulated in %.1f s" % (name, T, N, gen_s))
    world, rank = D.world, D.rank
    # --gpus N without torchrun: ONE process, the alignment sharded over N devices behind the C ABI (b200lk_create_sharded)
    ndev_sp = args.gpus if (world == 1 and args.gpus > 1) else 0
    b, e = MG.shard_range(N, world, rank) if world > 1 else (0, N)
    shard = aln if world == 1 else W.Alignment(aln.nstates, aln.names, np.ascontiguousarray(aln.states[:, b:e]),
                                               np.ascontiguousarray(aln.ptn_freq[b:e]))
    eng = engine_for(shard, model, rates, props, nslots=T, device=D.local, devices=list(range(ndev_sp)) if ndev_sp else None)
    root = tree.adj[0][0]
    tr = W.traversal(tree, root, 0)
    packed = pack_traversal(tr)
    dad, node, L = (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, root)
    tipl = W.tip_partial_lh(model)
    model_args = (model.eval, model.evec, model.inv_evec, np.asarray(rates), np.asarray(props), model.mixw, tipl)
    h2d = sum(np.asarray(a).nbytes for a in model_args) + 16 * packed[1] + 16 * packed[3]
    if ndev_sp:
        r = timed_sharded_run(eng, packed, dad, node, L, model_args, steps, warmup, ndev_sp)
    else:
        r = timed_engine_run(D, eng, packed, dad, node, L, model_args, steps, warmup, args.collective)
    gpu_plh = eng.fetch_pattern_lh()
    gpu_scale = eng.fetch_scale_num(tr[-1][2])
    eng.close()
    weak = None
    if world > 1 and weak_too:
        # secondary figure: the WHOLE alignment on every GPU (weak scaling).  Every rank holds the same data, so the
        # all-reduced sums must be world x one rank's own
        engw = engine_for(aln, model, rates, props, nslots=T, device=D.local)
        rw = timed_engine_run(D, engw, packed, dad, node, L, model_args, steps, warmup, args.collective)
        engw.close()
        one = rw["local_result"]
        if one is None:  # peer exchange: every rank holds the same data, so one rank's own sums are total / world;
            # checked against an evaluation of the same alignment without peers
            eng1 = engine_for(aln, model, rates, props, nslots=T, device=D.local)
            one = eng1.evaluate(tr, dad, node, L)
            eng1.close()
            rw["local_result"] = one
        weak = {"value": len(tr) * N * world * steps / (rw["ms"] * 1e-3), "unit": UNIT, "scaling": "weak",
                "patterns_per_gpu": N, "ms_per_step": rw["ms"] / steps, "kernel_ms_per_step": rw["partial_ms"] / steps,
                "e2e_value": len(tr) * N * world * steps / (rw["e2e_ms"] * 1e-3),
                "allreduce_check_rel": rel(np.asarray(rw["result"]), world * np.asarray(rw["local_result"]))}
        weak["ok"] = bool(weak["allreduce_check_rel"] <= TOL)
    if rank != 0:
        return None
    log("%s: GPU timed (%.3f ms per step, e2e %.3f ms)" % (name, r["ms"] / steps, r["e2e_ms"] / steps))
    ncm = ncat * nmix
    roof, kinds = roofline_of(n, ncm, tr, e - b, r, steps)
    if ndev_sp:  # bytes are summed over the shards, the kernel time is the slowest shard's: per-device bandwidth
        for obj in (roof, roof.get("hbm", {})):
            if obj.get("achieved") is not None and obj.get("unit") == "GB/s":
                obj["achieved"] /= ndev_sp
                obj["frac"] = obj["achieved"] / obj["peak"]
                obj["per_device"] = True
        if roof.get("unit") == "TFLOP/s" and roof.get("achieved") is not None:
            roof["achieved"] /= ndev_sp
            roof["frac"] = roof["achieved"] / roof["peak"]
            roof["per_device"] = True
    attach_ncu_traffic(roof, name, world == 1 and not ndev_sp and N == N0)
    updates_per_step = len(tr) * N  # strong scaling: the named shape is the total over all ranks
    line = {
        "metric": METRIC, "value": updates_per_step * steps / (r["ms"] * 1e-3), "unit": UNIT,
        "n_gpus": ndev_sp or world, "steps": steps, "warmup": max(3, warmup), "ms_per_step": r["ms"] / steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic (seeded AliSim-equivalent sampler, every column its own pattern; the same alignment on every rank)",
        "process_model": ("one process drives %d devices through b200lk_create_sharded; per-shard sums added on the host in "
                          "device-index order; timed by wall clock around blocking steps" % ndev_sp) if ndev_sp
        else "one process per GPU (torchrun), NCCL all-reduce of 3 doubles" if world > 1 else "one process, one GPU",
        "config": {"workload": name + ": " + WORKLOADS[name][7], "model": model_name, "ntaxa": T, "patterns": N,
                   "patterns_per_gpu": (e - b) // (ndev_sp or 1), "nstates": n, "ncat": ncat, "nmix": nmix,
                   "node_kinds_TT_TI_II": [kinds[0], kinds[1], kinds[2]],
                   "sharding": "contiguous pattern ranges (whole 32-pattern tiles) of ONE alignment, one per GPU; "
                               "NCCL all-reduce of [lnL, df, ddf] only",
                   "l2": "inputs larger than L2 (partials %.1f GB per GPU rewritten every step)" % (len(tr) * (e - b) * n * ncm * 8 / 1e9)},
        "evals_per_sec": steps / (r["ms"] * 1e-3),
        "e2e": {"value": updates_per_step * steps / (r["e2e_ms"] * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 24,
                "ms_per_step": r["e2e_ms"] / steps, "evals_per_sec": steps / (r["e2e_ms"] * 1e-3)},
        "gpu_launches": int(r["launches"]), "launches_per_step": r["launches"] / steps,
        "clocks": r["clocks"], "roofline": roof, "lnl_df_ddf": list(r["result"]),
    }
    if world > 1:
        line["collective"] = {
            "used": "in-kernel exchange over NVLink peer memory (b200lk_peer_connect): the last CTA of the finishing kernel "
                    "stores the 3 sums into every peer's buffer, waits for the peers' and adds in rank order" if not r.get("used_nccl")
            else "ncclAllReduce(sum) of 3 doubles on the evaluation's stream" + (" (peer exchange unavailable: %s)" % r["peer_note"] if r.get("peer_note") else ""),
            "nccl_allreduce_3_doubles_us": r["collective_us"]}
        line["weak_scaling"] = weak
    # ---- parity against the reference kernel on the same patterns, and the CPU baseline ----
    if not args.no_cpu:
        t0 = time.perf_counter()
        cpu = CpuLeg(tree, model, rates, props, aln, nptn=cpu_patterns or args.cpu_sample or None)
        if want_cpu_baseline:
            dt, cpu_res = cpu.timed(max(1, args.cpu_steps), 1)
            line["cpu_baseline"] = cpu.baseline(dt, max(1, args.cpu_steps), N)
        else:
            cpu_res = cpu.evaluate()
        if cpu.full:
            gres, gplh, gsc = r["result"], gpu_plh, gpu_scale
            extra = {"patterns": N, "same_patterns_as_timed_run": True}
        else:
            # the CPU leg holds a prefix: the GPU evaluates exactly that prefix, and the per-pattern values of the
            # FULL-size timed run are compared on it as well
            gres, gplh, gsc = gpu_eval_patterns(tree, model, rates, props, aln, cpu.nptn, D.local)
            extra = {"patterns": cpu.nptn, "same_patterns_as_timed_run": False,
                     "ram_limited": cpu.ram_limited,
                     "full_run_pattern_lh_rel": rel(gpu_plh[:min(len(gpu_plh), cpu.nptn)], cpu.pattern_lh()[:len(gpu_plh)])}
        if world == 1:
            extra["lnl_equals_sum_f_pattern_lh_rel"] = rel(r["result"][0], float(np.dot(aln.ptn_freq, gpu_plh)))
        line["parity"] = parity_block(gres, cpu_res, gplh, cpu.pattern_lh(), gsc, cpu.top_scale(), cpu.kind, extra)
        line["parity"]["seconds"] = time.perf_counter() - t0
        cpu.close()
        log("%s: parity %s (lnl %.2e df %.2e ddf %.2e per-pattern %.2e)" % (
            name, "ok" if line["parity"]["ok"] else "FAILED", line["parity"]["lnl_rel"], line["parity"]["df_rel"],
            line["parity"]["ddf_rel"], line["parity"]["pattern_lh_rel"]))
    if weak is not None and not weak["ok"]:
        line.setdefault("parity", {})["ok"] = False
    return line


def newick_of(tree):
    names = tree.names
    tree.names = ["T%d" % i for i in range(tree.ntaxa)]
    try:
        return tree.to_newick(0, 17)
    finally:
        tree.names = names


def measure_partitioned(args, D, steps, warmup):
    """cfg5: supermatrix of `--genes` genes x ~1k patterns, alternating GTR+G4 / LG+G4, edge-linked proportional
    branch lengths, through the C++ host layer's PhyloSuperTree (batched launches over the genes of equal dimensions).
    A step = every partial of every gene stale -> PhyloSuperTree::computeLikelihood.  N > 1: whole genes placed on
    GPUs by LPT on the reference's cost T*N*n (b200::placePartitions); the only exchange is the all-reduce of the
    summed lnL.  Parity: per-gene lnL against the reference's own PhyloTree on the first genes of this rank."""
    import torch
    from iqtree2_b200 import build_host, host as H, multigpu as MG, workload as W
    rank, world, local = D.rank, D.world, D.local
    build_host.build()
    T, G = 100, args.genes
    rng = np.random.default_rng(4)
    lengths = rng.integers(500, 1501, size=G)
    part_rates = rng.gamma(2.0, 0.5, size=G)
    nst = np.where(np.arange(G) % 2 == 0, 4, 20)
    cost = T * lengths.astype(float) * nst
    place = MG.place_partitions(cost, world)
    mine = [g for g in range(G) if place[g] == rank]
    tree = W.random_tree(T, 4)
    st = H.HostSuperTree(newick_of(tree), edge_linked=True, device=local)
    dna, dna_name = model_for(4, 1, 4)
    aa, aa_name = model_for(20, 1, 4)
    rates, props = W.discrete_gamma(0.5)
    updates, alg_bytes, min_bytes = 0, 0, 0
    tr0 = W.traversal(tree, tree.adj[0][0], 0)
    keep = {}
    for pos, g in enumerate(mine):
        model = dna if nst[g] == 4 else aa
        scaled = W.random_tree(T, 4)
        for (u, v) in scaled.branches():
            scaled.set_length(u, v, min(9.0, tree.get_length(u, v) * part_rates[g]))
        aln = W.simulate_alignment(scaled, model, rates, props, int(lengths[g]), 5000 + g, compress=False)
        st.add_partition("gene%d" % g, H.HostAlignment.from_states(aln.states, int(nst[g]), compress=False),
                         H.HostModel.from_eigen(model), H.HostRate.categories(rates, props), float(part_rates[g]))
        updates += (T - 2) * aln.nptn
        ab, _ = traversal_bytes(tr0, aln.nptn, int(nst[g]), 4)
        alg_bytes += ab
        # 4-state genes: the pattern-stationary traversal writes each vector once and reads tips; 20-state genes are
        # level-batched and move the algorithmic bytes
        min_bytes += aln.nptn * ((T - 2) * (8 * 16 + 1) + T) if nst[g] == 4 else ab
        if len(keep) < args.parity_genes:
            keep[pos] = (scaled, model, aln)  # part_scores() is in add_partition order
    st.set_batching(not args.no_batch)
    st.initialize()
    total = torch.zeros(1, dtype=torch.float64, device="cuda")

    def step():
        st.clear_partials()
        lnl = st.compute_likelihood()
        if D.dist:
            total[0] = lnl
            D.dist.all_reduce(total)
            return float(total[0])
        return lnl

    for _ in range(max(3, warmup)):
        lnl = step()
    calls0, launches0 = st.kernel_calls(), st.gpu_launches()
    sampler = ClockSampler(local)
    sampler.start()
    D.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    for _ in range(steps):
        lnl = step()
    e1.record()
    D.sync()
    wall_ms = (time.perf_counter() - t0) * 1e3
    ms = max(e0.elapsed_time(e1), wall_ms)  # every step ends with a blocking read: wall clock and events agree
    clocks = sampler.finish()
    scores = st.part_scores()
    # per-gene parity against the reference's own PhyloTree (the loop of tree/phylosupertree.cpp:701-725)
    parity = None
    if not args.no_cpu and keep:
        import oracle
        worst, ngenes = 0.0, 0
        for idx, (scaled, model, aln) in keep.items():
            if oracle.ref_available():
                ref = oracle.RefRun(scaled, aln, model, rates, props, kernel=oracle.LK_AVX_FMA, threads=host_cores()[0])
                want = ref.lnl_at_leaf(0)
                ref.close()
                kind = "reference"
            else:
                from iqtree2_b200.capi import INTERNAL, LEAF
                o = oracle.oracle_engine(aln, model, rates, props, nslots=T)
                trg = W.traversal(scaled, scaled.adj[0][0], 0)
                o.update_partials(trg)
                want = o.branch_lnl((LEAF, 0), (INTERNAL, trg[-1][2]), scaled.get_length(0, scaled.adj[0][0]))
                o.close()
                kind = "port"
            worst = max(worst, rel(scores[idx], want))
            ngenes += 1
        parity = {"against": kind, "genes_compared": ngenes, "per_gene_lnl_rel": worst, "tol": TOL,
                  "sum_of_gene_lnl_equals_total_rel": rel(float(np.sum(scores)), lnl) if world == 1 else None,
                  "ok": bool(worst <= TOL)}
    ms, = D.max(ms)
    import torch as _t
    agg = _t.tensor([float(updates), float(alg_bytes), float(min_bytes)], dtype=_t.float64, device="cuda")
    if D.dist:
        D.dist.all_reduce(agg)
    updates_all, alg_all, min_all = (float(x) for x in agg.cpu())
    if rank != 0:
        return None
    (hbm, hbm_src), _ = peaks()
    gbs = min_all * steps / (ms * 1e-3) / 1e9
    val = updates_all * steps / (ms * 1e-3)
    return {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": world, "steps": steps,
            "warmup": max(3, warmup), "ms_per_step": ms / steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic (seeded AliSim-equivalent sampler per gene)",
            "config": {"workload": "cfg5: partitioned supermatrix, %d genes x U(500,1500) patterns, 100 taxa, alternating "
                                   "%s+G4 / %s+G4, edge-linked proportional; lnL evaluation of the whole supermatrix" % (G, dna_name.split("{")[0], aa_name),
                       "genes": G, "genes_this_rank": len(mine), "placement": "LPT on T*N*n",
                       "l2": "every gene's partial vectors are rewritten each step (%.1f GB in all, larger than L2)" % (min_all / 1e9)},
            "evals_per_sec": steps / (ms * 1e-3),
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": int(G * (198 * 16 + 98 * 16)),
                    "d2h_bytes_per_step": 8 * G,
                    "note": "the timed call IS the host-facing call (b200::PhyloSuperTree::computeLikelihood): traversal "
                            "descriptors H2D and per-gene lnL D2H are inside the timed region"},
            "gpu_launches": int(st.gpu_launches() - launches0),
            "kernel_abi_calls_per_step": (st.kernel_calls() - calls0) / steps,
            "batches": dict(zip(("batches", "genes_in_batches", "batched_calls"), st.batch_info())),
            "clocks": clocks, "lnl": lnl, "parity": parity,
            "roofline": {"bound": "hbm", "kernel": "whole supermatrix evaluation (all launches of a step, wall clock): "
                                                   "launch- and level-bound regime of ~1k-pattern genes",
                         "achieved": gbs, "peak": hbm, "peak_source": hbm_src, "unit": "GB/s", "frac": gbs / hbm,
                         "traffic": None, "min_dram_bytes_per_step": int(min_all),
                         "algorithmic_bytes_per_step": int(alg_all)}}


def measure_mem_save(args, D, name, steps, warmup, mem_save, parity_patterns):
    """A workload whose partial vectors do not fit one GPU with one slot per internal node (cfg4: 98 x 3.84 GB at
    100k patterns), evaluated on ONE GPU through b200::PhyloTree with LM_MEM_SAVE (-mem, tree/memslot.cpp): the
    traversal builder recycles log2(T)+1 slots.  A step = every partial stale -> computeLikelihood."""
    import torch
    from iqtree2_b200 import build_host, host as H, workload as W
    build_host.build()
    T, N, n, ncat, nmix, *_ = WORKLOADS[name]
    tree, model, rates, props, aln, model_name = build_workload(name, args.patterns or None)
    N = aln.nptn
    ha = H.HostAlignment.from_states(aln.states, n, compress=False)
    t = H.HostTree(newick_of(tree), ha, H.HostModel.from_eigen(model), H.HostRate.categories(rates, props),
                   mem_save=mem_save, device=D.local)
    slots = t.max_lh_slots
    updates = (T - 2) * N

    def step():
        t.clear_partials()
        return t.compute_likelihood()

    for _ in range(max(3, warmup)):
        lnl = step()
    sampler = ClockSampler(D.local)
    sampler.start()
    torch.cuda.synchronize()
    launches0 = t.gpu_launches()
    t0 = time.perf_counter()
    for _ in range(steps):
        lnl = step()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.finish()
    launches = int(t.gpu_launches() - launches0)
    lnl, plh = t.compute_likelihood(want_pattern_lh=True)
    t.close()
    block = n * ncat * nmix
    tr = W.traversal(tree, tree.adj[0][0], 0)
    flops = traversal_flops(tr, N, n, ncat * nmix)
    abytes, _ = traversal_bytes(tr, N, n, ncat * nmix)
    (hbm, hbm_src), (f64, f64_src) = peaks()
    tfl = flops * steps / (ms * 1e-3) / 1e12
    gbs = abytes * steps / (ms * 1e-3) / 1e9
    val = updates * steps / (ms * 1e-3)
    line = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": 1, "steps": steps, "warmup": max(3, warmup),
            "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic (seeded AliSim-equivalent sampler, every column its own pattern)",
            "config": {"workload": "%s through b200::PhyloTree with -mem: %s" % (name, WORKLOADS[name][7]),
                       "model": model_name, "ntaxa": T, "patterns_per_gpu": int(N), "nstates": n, "ncat": ncat, "nmix": nmix,
                       "lh_slots": slots, "slots_without_mem_save": T - 2,
                       "partial_vectors_gb": slots * N * block * 8 / 1e9,
                       "partial_vectors_gb_without_mem_save": (T - 2) * N * block * 8 / 1e9,
                       "l2": "every step rewrites %.1f GB of partial vectors" % (updates * block * 8 / 1e9)},
            "evals_per_sec": steps / (ms * 1e-3),
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": int((T - 2) * 16 + (2 * T - 3) * 16), "d2h_bytes_per_step": 8,
                    "note": "the timed call IS the host-facing call (b200::PhyloTree::computeLikelihood)"},
            "gpu_launches": launches, "clocks": clocks, "lnl": lnl,
            "roofline": {"bound": "tensor" if n >= 9 else "hbm",
                         "kernel": "whole evaluation through the host layer (wall clock; every launch of a step, slot "
                                   "recycling included): algorithmic flops of the T-2 updates / step time",
                         "achieved": tfl, "peak": f64, "peak_source": f64_src, "unit": "TFLOP/s", "frac": tfl / f64,
                         "traffic": None, "algorithmic_flops_per_step": int(flops),
                         "hbm": {"achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                                 "algorithmic_bytes_per_step": int(abytes)}}}
    if not args.no_cpu and parity_patterns:
        cpu = CpuLeg(tree, model, rates, props, aln, nptn=parity_patterns)
        cpu_res = cpu.evaluate()
        gres, gplh, gsc = gpu_eval_patterns(tree, model, rates, props, aln, cpu.nptn, D.local)
        extra = {"patterns": cpu.nptn, "same_patterns_as_timed_run": False,
                 "full_run_pattern_lh_rel": rel(plh[:cpu.nptn], cpu.pattern_lh()),
                 "full_run_lnl_equals_sum_f_pattern_lh_rel": rel(lnl, float(np.dot(aln.ptn_freq, plh)))}
        line["parity"] = parity_block(gres, cpu_res, gplh, cpu.pattern_lh(), gsc, cpu.top_scale(), cpu.kind, extra)
        cpu.close()
    return line


def measure_sweep(args, D, name, sweeps, cpu_patterns):
    """One `optimizeAllBranches` sweep (tree/phylotree.cpp:2689-2765 -> optimizeOneBranch :2585-2635 -> minimizeNewton,
    utils/optimization.cpp:422-499) through b200::PhyloTree: the caller that issues most of the path's calls in a
    real run.  Checked against the reference's own optimizeAllBranches on a pattern prefix (lnL after the sweep)."""
    import torch
    from iqtree2_b200 import build_host, host as H, workload as W
    build_host.build()
    T, N, n, *_ = WORKLOADS[name]
    tree, model, rates, props, aln, model_name = build_workload(name, args.patterns or None)
    ha = H.HostAlignment.from_states(aln.states, n, compress=False)
    t = H.HostTree(newick_of(tree), ha, H.HostModel.from_eigen(model), H.HostRate.categories(rates, props), device=D.local)
    start = t.lengths()
    lnl0 = t.compute_likelihood()
    t.optimize_all_branches(iterations=1)  # warm-up: allocations, first launches
    sampler = ClockSampler(D.local)
    sampler.start()
    modes = {}
    for mode, on in (("device_resident_lengths", True), ("branch_by_branch", False)):
        t.set_device_sweep(on)
        runs = []
        for _ in range(sweeps):
            t.set_lengths(start)
            t.clear_partials()
            t.compute_likelihood()
            torch.cuda.synchronize()
            l0, k0, d0 = t.gpu_launches(), t.kernel_calls(), t.device_sweeps()
            t0 = time.perf_counter()
            lnl = t.optimize_all_branches(iterations=1)
            torch.cuda.synchronize()
            runs.append(((time.perf_counter() - t0) * 1e3, lnl, t.gpu_launches() - l0, t.kernel_calls() - k0, t.device_sweeps() - d0))
        modes[mode] = {"ms_per_sweep": min(r[0] for r in runs), "lnl_after_one_sweep": runs[-1][1], "launches_per_sweep": int(runs[-1][2]),
                       "abi_calls_per_sweep": int(runs[-1][3]), "ran_on_device": bool(runs[-1][4])}
    t.set_device_sweep(True)
    clocks = sampler.finish()
    nb = int(t.nbranches)
    best = modes["device_resident_lengths"]["ms_per_sweep"]
    out = {"workload": name, "model": model_name, "ntaxa": T, "patterns": int(aln.nptn), "branches": nb, "lnl_start": lnl0,
           "ms_per_sweep": best, "us_per_branch": best * 1e3 / nb, "launches_per_sweep": modes["device_resident_lengths"]["launches_per_sweep"],
           "lnl_after_one_sweep": modes["device_resident_lengths"]["lnl_after_one_sweep"], "modes": modes,
           "modes_agree_rel": rel(modes["device_resident_lengths"]["lnl_after_one_sweep"], modes["branch_by_branch"]["lnl_after_one_sweep"]),
           "pattern_branches_per_sec": nb * aln.nptn / (best * 1e-3), "clocks": clocks}
    t.close()
    if not args.no_cpu and cpu_patterns:
        import oracle
        if oracle.ref_available():
            M = min(cpu_patterns, aln.nptn)
            sub = W.Alignment(aln.nstates, aln.names, np.ascontiguousarray(aln.states[:, :M]), np.ascontiguousarray(aln.ptn_freq[:M]))
            ref = oracle.RefRun(tree, sub, model, rates, props, kernel=oracle.LK_AVX_FMA, threads=host_cores()[0])
            ref.compute_likelihood()
            t0 = time.perf_counter()
            rl = ref.optimize_all_branches(iterations=1)
            dt = time.perf_counter() - t0
            ref.close()
            # the same prefix through the GPU host layer: lnL after one sweep must agree with the reference's
            hs = H.HostAlignment.from_states(sub.states, n, compress=False)
            ts = H.HostTree(newick_of(tree), hs, H.HostModel.from_eigen(model), H.HostRate.categories(rates, props), device=D.local)
            ts.compute_likelihood()
            gl = ts.optimize_all_branches(iterations=1)
            ts.close()
            out["cpu"] = {"kind": "reference", "cores": host_cores()[0], "patterns": M, "ms_per_sweep": dt * 1e3,
                          "lnl_after_one_sweep": rl, "pattern_branches_per_sec": nb * M / dt}
            out["parity"] = {"against": "reference optimizeAllBranches, one sweep, %d patterns" % M,
                             "lnl_after_sweep_rel": rel(gl, rl), "tol": 1e-7, "ok": bool(rel(gl, rl) <= 1e-7)}
    return out


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on this box's host cores, on the same
    config, metric and unit; each step one full evaluation of the full workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = args.workload
    T, N, n, ncat, nmix, *_ = WORKLOADS[name]
    tree, model, rates, props, aln, model_name = build_workload(name, args.patterns or None)
    cpu = CpuLeg(tree, model, rates, props, aln, nptn=args.cpu_sample or None)
    steps = max(1, args.steps)
    # keep the whole run within a few minutes: one evaluation is timed first, the step count shrinks if needed
    dt1, _ = cpu.timed(1, max(1, min(args.warmup, 2)))
    budget = 150.0
    steps_run = max(1, min(steps, int(budget / max(dt1, 1e-6))))
    dt, res = cpu.timed(steps_run, 0)
    b = cpu.baseline(dt, steps_run, aln.nptn)
    line = {"impl": "reference", "metric": METRIC, "value": b["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "steps_timed": steps_run, "warmup": args.warmup, "ms_per_step": b["ms_per_eval"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic (seeded AliSim-equivalent sampler, every column its own pattern)",
            "config": {"workload": name + ": " + WORKLOADS[name][7], "model": model_name, "ntaxa": T, "patterns": int(aln.nptn),
                       "patterns_timed": int(cpu.nptn), "nstates": n, "ncat": ncat, "nmix": nmix},
            "cpu_baseline": {k: b[k] for k in ("value", "unit", "cores", "kind", "sample", "physical_cores", "logical_cpus",
                                               "pinning", "same_config")},
            "evals_per_sec": b["evals_per_sec"], "gpu_launches": 0, "lnl_df_ddf": [float(x) for x in res],
            "e2e": {"value": b["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def shape_summary(line):
    """what a sub-result keeps of a full line"""
    keep = ("value", "unit", "ms_per_step", "evals_per_sec", "n_gpus", "steps", "gpu_launches", "clocks", "roofline", "parity",
            "lnl", "lnl_df_ddf", "batches")
    out = {k: line[k] for k in keep if k in line}
    out["workload"] = line["config"]["workload"]
    out["model"] = line["config"].get("model")
    out["e2e_value"] = line["e2e"]["value"]
    out["frac"] = line["roofline"].get("frac")
    out["peak_source"] = line["roofline"].get("peak_source")
    return out


def main():
    own_stdout()
    host_cores()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--mem-save", type=float, default=None, metavar="FRACTION",
                    help="run the workload through the host layer with -mem slot recycling (0: log2(T)+1 slots)")
    ap.add_argument("--no-batch", action="store_true", help="cfg5: one launch sequence per gene instead of batched launches")
    ap.add_argument("--patterns", type=int, default=0, help="override the number of patterns (debug)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="patterns in the CPU legs (0 = the full workload)")
    ap.add_argument("--cpu-steps", type=int, default=3, help="evaluations timed for cpu_baseline")
    ap.add_argument("--no-cpu", action="store_true", help="skip the parity and cpu_baseline legs")
    ap.add_argument("--collective", default="peer", choices=["peer", "nccl"],
                    help="N > 1: how [lnL, df, ddf] are summed over the ranks (in-kernel peer exchange, or ncclAllReduce)")
    ap.add_argument("--genes", type=int, default=1000, help="cfg5: number of genes")
    ap.add_argument("--parity-genes", type=int, default=8, help="cfg5: genes compared with the reference")
    ap.add_argument("--shapes", default="auto", help="sub-results for the other BASELINE shapes in the same line: "
                    "auto (cfg2 at N=1: all), none, or a comma list of cfg3,cfg4_full_mem,cfg5,branch_sweep_cfg2")
    ap.add_argument("--time-budget", type=float, default=420.0, help="seconds after which remaining shapes are skipped")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    D = Dist()
    steps, warmup = max(1, args.steps), args.warmup
    if args.workload == "cfg5":
        line = measure_partitioned(args, D, steps, warmup)
    elif args.mem_save is not None:
        line = measure_mem_save(args, D, args.workload, steps, warmup, args.mem_save, 2000)
    else:
        line = measure_alignment(args, D, args.workload, steps, warmup, want_cpu_baseline=(D.world == 1))
    ok = True
    if D.rank == 0:
        want = args.shapes
        if want == "auto":
            want = "branch_sweep_cfg2,cfg3,cfg4_full_mem,cfg5" if (args.workload == "cfg2" and D.world == 1 and args.gpus == 1
                                                                   and not args.patterns and args.mem_save is None) else "none"
        shapes = {}
        for sh in [x for x in want.split(",") if x and x != "none"]:
            if time.perf_counter() - T_START > args.time_budget:
                shapes[sh] = {"skipped": "time budget of %.0f s reached" % args.time_budget}
                continue
            log("shape %s ..." % sh)
            try:
                if sh == "branch_sweep_cfg2":
                    shapes[sh] = measure_sweep(args, D, "cfg2", 3, 100_000)
                elif sh == "cfg3":
                    _WL_CACHE.clear()
                    shapes[sh] = shape_summary(measure_alignment(args, D, "cfg3", 5, 3, want_cpu_baseline=False, cpu_patterns=20_000))
                elif sh == "cfg4_full_mem":
                    _WL_CACHE.clear()
                    shapes[sh] = shape_summary(measure_mem_save(args, D, "cfg4full", 3, 3, 0.0, 2000))
                elif sh == "cfg5":
                    _WL_CACHE.clear()
                    shapes[sh] = shape_summary(measure_partitioned(args, D, 5, 3))
                else:
                    shapes[sh] = {"skipped": "unknown shape"}
            except Exception as ex:  # a sub-result never takes the headline down; it is reported as failed
                shapes[sh] = {"failed": "%s: %s" % (type(ex).__name__, ex)}
                ok = False
            log("shape %s done" % sh)
        if shapes:
            line["shapes"] = shapes
        for obj in [line] + [v for v in shapes.values() if isinstance(v, dict)]:
            par = obj.get("parity")
            if par is not None and not par.get("ok", True):
                ok = False
        line["parity_ok"] = ok
        line["bench_seconds"] = time.perf_counter() - T_START
        emit(line)
    D.close()
    if not ok:
        sys.stderr.write("bench.py: PARITY FAILED or a sub-result failed (see the parity blocks of the JSON line)\n")
        sys.exit(3)


if __name__ == "__main__":
    main()
