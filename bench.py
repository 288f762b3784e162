#!/usr/bin/env python
"""bench.py -- the tree-likelihood hot path on B200, one JSON line per run.

A *step* is one full evaluation of the workload through the C ABI (include/b200lk.h):
every partial-likelihood vector stale -> post-order update of all T-2 internal nodes ->
lnL at the root branch -> theta -> (df, ddf): SURVEY.md section 8(d), metric (1)+(2).

  value  pattern-node likelihood updates/s, device-timed (CUDA events on the launching stream),
         inputs resident in HBM, results left on the device (b200lk_evaluate_async)
  e2e    the same through the blocking call a host caller makes per model probe:
         b200lk_set_model (eigen-system + rates H2D) + b200lk_evaluate (traversal descriptors and
         branch lengths H2D, lnL/df/ddf D2H)
  roofline      the partial-update kernel: algorithmic bytes (B_K per node kind) /
                its CUDA-event time inside the timed region, against MEASURED_PEAKS.json
  cpu_baseline  the CPU checker (oracle/) on a bounded pattern sample, all host threads

N > 1 (torchrun): every rank owns an independent pattern shard of the same size (weak scaling); the
only exchange is an NCCL all-reduce of [lnL, df, ddf] in place on the device result buffer.

`--impl reference` times the CPU implementation only (oracle/_ref when built, else the oracle port).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "pattern_node_likelihood_updates_per_sec"
UNIT = "updates/s"

WORKLOADS = {
    # name: (ntaxa, npatterns, nstates, ncat, nmix, alpha, seed, description)
    "cfg2": (200, 1_000_000, 4, 4, 1, 0.5, 1, "synthetic DNA 200 taxa x 1M patterns, GTR+G4, full traversal + lnL + derv"),
    "cfg3": (500, 200_000, 20, 4, 1, 0.7, 2, "synthetic protein 500 taxa x 200k patterns, 20-state reversible +G4"),
    # cfg4 is 100k patterns over 8 GPUs (376 GB of partial vectors in all): 12 500 patterns = 47 GB per GPU
    "cfg4": (100, 12_500, 20, 4, 60, 0.7, 3, "synthetic protein 100 taxa x 100k patterns / 8 GPUs (12.5k per GPU), 20-state 60-class mixture +G4"),
    "cfg5": (100, 1000, 4, 4, 1, 0.5, 4, "partitioned supermatrix (see run_partitioned)"),
    "small": (50, 100_000, 4, 4, 1, 0.5, 5, "synthetic DNA 50 taxa x 100k patterns, GTR+G4 (smoke size)"),
    # not a BASELINE config: the 61-state codon case north_star names for the tensor path
    "codon": (100, 50_000, 61, 4, 1, 0.5, 6, "synthetic codon 100 taxa x 50k patterns, 61-state reversible +G4"),
}


def traversal_flops(tr, N, nstates, ncat_mix):
    """Algorithmic flops of the partial updates of one traversal:
    F_K = C * (2 n^2 (k_int + 1) + n) per pattern for a bifurcating node with k_int internal children."""
    total = 0
    for e in tr:
        k = sum(1 for ch in e[3] if ch[0] == 1)
        total += ncat_mix * (2 * nstates * nstates * (k + 1) + nstates)
    return total * N


def build_workload(name, npatterns=None, seed_shift=0):
    from iqtree2_b200 import workload as W
    T, N, n, ncat, nmix, alpha, seed, _ = WORKLOADS[name]
    if npatterns:
        N = npatterns
    tree = W.random_tree(T, seed)
    model = W.gtr_model() if (n == 4 and nmix == 1) else W.random_reversible_model(n, seed + 100, nmix)
    rates, props = W.discrete_gamma(alpha, ncat)
    aln = W.simulate_alignment(tree, model, rates, props, N, seed + 1000 + seed_shift, compress=False)
    return tree, model, rates, props, aln


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

    def summary(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def time_cpu(tree, model, rates, props, aln, sample, steps, warmup):
    """CPU baseline on the first `sample` patterns of the same alignment, all host threads: the reference's
    own AVX+FMA kernel with OpenMP (oracle/_ref, kind "reference") when that library was built, else the C
    restatement (kind "port").  One evaluation = all partials stale -> lnL at the root branch -> theta ->
    df/ddf, exactly the GPU step."""
    import oracle
    from iqtree2_b200 import workload as W
    from iqtree2_b200.capi import INTERNAL, LEAF
    sub = W.Alignment(aln.nstates, aln.names, np.ascontiguousarray(aln.states[:, :sample]),
                      np.ascontiguousarray(aln.ptn_freq[:sample]))
    cores = os.cpu_count() or 1
    c = tree.adj[0][0]
    tr = W.traversal(tree, c, 0)
    if oracle.ref_available():
        kind = "reference"
        ref = oracle.RefRun(tree, sub, model, rates, props, kernel=oracle.LK_AVX_FMA, threads=cores)

        def one():
            lnl = ref.lnl_at_leaf(0)
            df, ddf = ref.derv()
            return lnl, df, ddf
    else:
        kind = "port"
        eng = oracle.oracle_engine(sub, model, rates, props, nslots=tree.ntaxa)
        eng._f("set_threads")(eng.ctx, cores)
        args = (tr, (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, c))

        def one():
            return eng.evaluate(*args)
    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        res = one()
    dt = (time.perf_counter() - t0) / steps
    what = "reference AVX+FMA kernel (-lk FMA), OpenMP" if kind == "reference" else "C restatement, OpenMP"
    return {"value": len(tr) * sub.nptn / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": "%s; first %d of %d patterns, same tree/model, %d full evaluations" % (what, sub.nptn, aln.nptn, steps),
            "ms_per_eval": dt * 1e3, "evals_per_sec": 1.0 / dt, "lnl": res[0]}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    T, N, n, ncat, nmix, *_ = WORKLOADS[args.workload]
    sample = min(args.cpu_sample, args.patterns or N)
    tree, model, rates, props, aln = build_workload(args.workload, sample)
    r = time_cpu(tree, model, rates, props, aln, sample, max(1, args.steps), max(1, args.warmup))
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_eval"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic (seeded AliSim-equivalent sampler)",
            "config": {"workload": args.workload + ": " + WORKLOADS[args.workload][7], "ntaxa": T,
                       "patterns_timed": sample, "nstates": n, "ncat": ncat},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "evals_per_sec": r["evals_per_sec"], "gpu_launches": 0,
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def run_partitioned(args):
    """cfg5: supermatrix of `--genes` genes x ~1k patterns, alternating GTR+G4 / 20-state+G4, edge-linked
    proportional branch lengths, through the C++ host layer's PhyloSuperTree (one context and stream per gene,
    all genes of an evaluation enqueued before any result is read).  A step = every partial of every gene stale
    -> PhyloSuperTree::computeLikelihood.  N > 1: whole genes placed on GPUs by LPT on the reference's cost
    T*N*n (b200::placePartitions); the only exchange is the all-reduce of the summed lnL."""
    import torch
    import torch.distributed as dist
    from iqtree2_b200 import build_host, host as H, multigpu as MG, workload as W
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback on this path")
    torch.cuda.set_device(local)
    if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
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
    names = tree.names
    tree.names = ["T%d" % i for i in range(T)]
    newick = tree.to_newick(0, 17)
    tree.names = names
    st = H.HostSuperTree(newick, edge_linked=True, device=local)
    dna, aa = W.gtr_model(), W.random_reversible_model(20, 104)
    rates, props = W.discrete_gamma(0.5)
    updates = 0
    for g in mine:
        model = dna if nst[g] == 4 else aa
        scaled = W.random_tree(T, 4)
        for (u, v) in scaled.branches():
            scaled.set_length(u, v, min(9.0, tree.get_length(u, v) * part_rates[g]))
        aln = W.simulate_alignment(scaled, model, rates, props, int(lengths[g]), 5000 + g, compress=False)
        st.add_partition("gene%d" % g, H.HostAlignment.from_states(aln.states, int(nst[g]), compress=False),
                         H.HostModel.from_eigen(model), H.HostRate.categories(rates, props), float(part_rates[g]))
        updates += (T - 2) * aln.nptn
    st.set_batching(not args.no_batch)
    st.initialize()
    total = torch.zeros(1, dtype=torch.float64, device="cuda")

    def step():
        st.clear_partials()
        lnl = st.compute_likelihood()
        if world > 1:
            total[0] = lnl
            dist.all_reduce(total)
            return float(total[0])
        return lnl

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        lnl = step()
    calls0 = st.kernel_calls()
    launches0 = st.gpu_launches()
    sampler = ClockSampler(local)
    sampler.start()
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        lnl = step()
    e1.record()
    sync_all()
    wall_ms = (time.perf_counter() - t0) * 1e3
    ms = max(e0.elapsed_time(e1), wall_ms)  # every step ends with a blocking read: wall clock and events agree
    sampler.stop_flag = True
    sampler.join()
    t = torch.tensor([ms, float(updates)], dtype=torch.float64, device="cuda")
    if world > 1:
        tm = t.clone()
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(t)
        ms, updates_all = float(tm[0]), float(t[1])
    else:
        updates_all = float(updates)
    if rank == 0:
        val = updates_all * args.steps / (ms * 1e-3)
        line = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic (seeded AliSim-equivalent sampler per gene)",
                "config": {"workload": "cfg5: partitioned supermatrix, %d genes x U(500,1500) patterns, 100 taxa, alternating "
                                       "GTR+G4 / 20-state+G4, edge-linked proportional; lnL evaluation of the whole supermatrix" % G,
                           "genes": G, "genes_this_rank": len(mine), "placement": "LPT on T*N*n",
                           "l2": "every gene's partial vectors are rewritten each step (%.1f GB in all, larger than L2)" % (updates_all * 8 * 48 / 1e9)},
                "evals_per_sec": args.steps / (ms * 1e-3),
                "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": int(G * (198 * 16 + 98 * 16)),
                        "d2h_bytes_per_step": 8 * G,
                        "note": "the timed call IS the host-facing call (b200::PhyloSuperTree::computeLikelihood): traversal "
                                "descriptors H2D and per-gene lnL D2H are inside the timed region"},
                "gpu_launches": int(st.gpu_launches() - launches0),
                "kernel_abi_calls_per_step": (st.kernel_calls() - calls0) / args.steps,
                "batches": dict(zip(("batches", "genes_in_batches", "batched_calls"), st.batch_info())),
                "clocks": sampler.summary(), "lnl": lnl,
                "roofline": {"bound": "hbm", "kernel": "launch-bound regime (1000 small contexts): see DESIGN.md 6",
                             "achieved": None, "peak": None, "unit": "GB/s", "frac": None, "traffic": None}}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_mem_save(args):
    """A workload whose partial vectors do not fit one GPU with one slot per internal node (cfg4: 98 x 3.84 GB at
    100k patterns), evaluated on ONE GPU through b200::PhyloTree with LM_MEM_SAVE (-mem, tree/memslot.cpp): the
    traversal builder recycles log2(T)+1 slots.  A step = every partial stale -> computeLikelihood."""
    import torch
    from iqtree2_b200 import build_host, host as H, workload as W
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback on this path")
    build_host.build()
    T, N, n, ncat, nmix, *_ = WORKLOADS[args.workload]
    N = args.patterns or N
    tree, model, rates, props, aln = build_workload(args.workload, N)
    names = tree.names
    tree.names = ["T%d" % i for i in range(T)]
    newick = tree.to_newick(0, 17)
    tree.names = names
    ha = H.HostAlignment.from_states(aln.states, n, compress=False)
    t = H.HostTree(newick, ha, H.HostModel.from_eigen(model), H.HostRate.categories(rates, props), mem_save=args.mem_save)
    slots = t.max_lh_slots
    updates = (T - 2) * aln.nptn

    def step():
        t.clear_partials()
        return t.compute_likelihood()

    for _ in range(max(3, args.warmup)):
        lnl = step()
    sampler = ClockSampler(0)
    sampler.start()
    torch.cuda.synchronize()
    launches0 = t.gpu_launches()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        lnl = step()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3
    sampler.stop_flag = True
    sampler.join()
    block = n * ncat * nmix
    val = updates * args.steps / (ms * 1e-3)
    line = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic (seeded AliSim-equivalent sampler, every column its own pattern)",
            "config": {"workload": "%s through b200::PhyloTree with -mem: %s" % (args.workload, WORKLOADS[args.workload][7]),
                       "ntaxa": T, "patterns_per_gpu": int(aln.nptn), "nstates": n, "ncat": ncat, "nmix": nmix,
                       "lh_slots": slots, "slots_without_mem_save": T - 2,
                       "partial_vectors_gb": slots * aln.nptn * block * 8 / 1e9,
                       "partial_vectors_gb_without_mem_save": (T - 2) * aln.nptn * block * 8 / 1e9,
                       "l2": "every step rewrites %.1f GB of partial vectors" % (updates * block * 8 / 1e9)},
            "evals_per_sec": args.steps / (ms * 1e-3),
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": int((T - 2) * 16 + (2 * T - 3) * 16), "d2h_bytes_per_step": 8,
                    "note": "the timed call IS the host-facing call (b200::PhyloTree::computeLikelihood)"},
            "gpu_launches": int(t.gpu_launches() - launches0), "clocks": sampler.summary(), "lnl": lnl,
            "roofline": {"bound": "tensor" if n >= 9 else "hbm", "kernel": "whole evaluation (wall clock)",
                         "achieved": None, "peak": None, "unit": "TFLOP/s", "frac": None, "traffic": None}}
    emit(line)


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


def main():
    own_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--mem-save", type=float, default=None, metavar="FRACTION",
                    help="run the workload through the host layer with -mem slot recycling (0: log2(T)+1 slots)")
    ap.add_argument("--no-batch", action="store_true", help="cfg5: one launch sequence per gene instead of batched launches")
    ap.add_argument("--patterns", type=int, default=0, help="override patterns per GPU (debug)")
    ap.add_argument("--cpu-sample", type=int, default=100_000, help="patterns in the CPU-baseline sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--genes", type=int, default=1000, help="cfg5: number of genes")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "cfg5":
        return run_partitioned(args)
    if args.mem_save is not None:
        return run_mem_save(args)

    import torch
    import torch.distributed as dist
    from iqtree2_b200 import workload as W
    from iqtree2_b200.capi import INTERNAL, LEAF, engine_for, pack_traversal

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback on this path")
    torch.cuda.set_device(local)
    if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"  # NCCL's version banner goes to stdout: keep stdout to the one JSON line
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    T, N, n, ncat, nmix, *_ = WORKLOADS[args.workload]
    if args.patterns:
        N = args.patterns
    # every rank: same tree and model, its own pattern shard (seed shifted by rank)
    tree, model, rates, props, aln = build_workload(args.workload, N, seed_shift=rank)
    eng = engine_for(aln, model, rates, props, nslots=T, device=local)
    # a dedicated (non-default) stream shared by the engine, NCCL and the timing events
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    eng.set_stream(stream.cuda_stream)
    root = tree.adj[0][0]
    tr = W.traversal(tree, root, 0)
    packed = pack_traversal(tr)
    dad, node, L = (LEAF, 0), (INTERNAL, tr[-1][2]), tree.get_length(0, root)
    res_t = eng.result_tensor()
    tipl = W.tip_partial_lh(model)
    model_args = (model.eval, model.evec, model.inv_evec, np.asarray(rates), np.asarray(props), model.mixw, tipl)
    h2d = sum(np.asarray(a).nbytes for a in model_args) + 16 * packed[1] + 16 * packed[3]
    alg_bytes, kinds = traversal_bytes(tr, N, n, ncat * nmix)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        eng.evaluate_async(packed, dad, node, L)
        if world > 1:
            dist.all_reduce(res_t)

    def step_e2e():
        eng.set_model(*model_args)
        eng.evaluate_async(packed, dad, node, L)
        if world > 1:
            dist.all_reduce(res_t)
        return eng.read_results()

    for _ in range(max(3, args.warmup)):
        step_device()
    sync_all()
    eng.synchronize()
    eng.get_profile(reset=True)
    eng.reset_counters()

    # ---- device-timed region -------------------------------------------------------------------
    sampler = ClockSampler(local)
    sampler.start()
    eng.set_profiling(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    sync_all()
    ms = e0.elapsed_time(e1)
    eng.set_profiling(False)
    partial_ms, partial_launches = eng.get_profile(reset=True)
    _, launches = eng.counters()
    result = eng.read_results()

    # ---- end-to-end region ---------------------------------------------------------------------
    for _ in range(2):
        step_e2e()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        result = step_e2e()
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    sampler.stop_flag = True
    sampler.join()

    t = torch.tensor([ms, e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = (float(x) for x in t.cpu())
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    updates_per_step = len(tr) * N * world
    peaks = {}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg_bytes * args.steps / (partial_ms * 1e-3) / 1e9 if partial_ms > 0 else None
    alg_flops = traversal_flops(tr, N, n, ncat * nmix)
    fp64_peak, fp64_src = 37.1, "fallback: DMMA peak measured once with tools/fp64_peak.cu (37.1)"
    fp = os.path.join(ROOT, "profiles", "r1g_fp64_peak.json")
    if os.path.exists(fp):
        fp64_peak, fp64_src = float(json.load(open(fp))["dmma_tflops"]), "profiles/r1g_fp64_peak.json (measured DMMA m8n8k4 peak, tools/fp64_peak.cu)"
    tflops = alg_flops * args.steps / (partial_ms * 1e-3) / 1e12 if partial_ms > 0 else None
    line = {
        "metric": METRIC, "value": updates_per_step * args.steps / (ms * 1e-3), "unit": UNIT,
        "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic (seeded AliSim-equivalent sampler, every column its own pattern)",
        "config": {"workload": args.workload + ": " + WORKLOADS[args.workload][7], "ntaxa": T,
                   "patterns_per_gpu": N, "nstates": n, "ncat": ncat, "nmix": nmix,
                   "node_kinds_TT_TI_II": [kinds[0], kinds[1], kinds[2]],
                   "sharding": "pattern shards, one per GPU; NCCL all-reduce of [lnL, df, ddf] only",
                   "l2": "inputs larger than L2 (partials %.1f GB per GPU)" % (len(tr) * N * n * ncat * nmix * 8 / 1e9)},
        "evals_per_sec": args.steps / (ms * 1e-3),
        "e2e": {"value": updates_per_step * args.steps / (e2e_ms * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 24,
                "ms_per_step": e2e_ms / args.steps, "evals_per_sec": args.steps / (e2e_ms * 1e-3)},
        "gpu_launches": int(launches),
        "clocks": sampler.summary(),
        "roofline": {"bound": "hbm", "kernel": "s4_traverse_kernel (4-state partial update, whole traversal)" if n == 4 else "partial update (pruning step)",
                     "achieved": achieved,
                     "peak": peak_gbs, "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                     "unit": "GB/s", "frac": (achieved / peak_gbs) if achieved else None, "traffic": None,
                     "algorithmic_bytes_per_step": int(alg_bytes), "kernel_ms_per_step": partial_ms / args.steps,
                     "launches_per_step": partial_launches / max(1, args.steps),
                     "note": "achieved = ALGORITHMIC bytes / kernel time; the 4-state kernel keeps children in registers/shared memory "
                             "across the traversal and moves fewer bytes than that (see traffic), so frac can exceed 1"},
        "lnl_df_ddf": list(result),
    }
    if not args.no_cpu and world == 1:
        sample = min(args.cpu_sample, N)
        line["cpu_baseline"] = {k: v for k, v in time_cpu(tree, model, rates, props, aln, sample, 3, 1).items()
                                if k in ("value", "unit", "cores", "kind", "sample", "evals_per_sec")}
    if n >= 9:
        # dense paths: FP64 tensor pipe (DMMA) is the bounding unit next to HBM (20 states: AI ~ 5 flop/B, at the
        # ridge; 61 states: ~ 15 flop/B, compute-bound)
        hbm = dict(line["roofline"])
        line["roofline"] = {"bound": "tensor", "kernel": "dense20_rp_kernel + dense20_tt_kernel (20-state partial update, FP64 DMMA, level-batched)" if n == 20
                            else "denseN_kernel (%d-state partial update, FP64 DMMA on 8-state tiles, level-batched)" % n,
                            "achieved": tflops, "peak": fp64_peak, "peak_source": fp64_src, "unit": "TFLOP/s",
                            "frac": tflops / fp64_peak if tflops else None, "traffic": None,
                            "algorithmic_flops_per_step": int(alg_flops), "kernel_ms_per_step": partial_ms / args.steps,
                            "launches_per_step": partial_launches / max(1, args.steps),
                            "hbm": {k: hbm[k] for k in ("achieved", "peak", "unit", "frac", "algorithmic_bytes_per_step")}}
    tr_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr_file):  # dram bytes per launch from the committed ncu --set full capture
        try:
            line["roofline"]["traffic"] = json.load(open(tr_file)).get(args.workload)
        except Exception:
            pass
    # physical side of the same kernel: measured DRAM bytes per launch / live launch duration, against the measured
    # stream ceilings of this pool's B200 (tools/hbm_streams.cu; the 4-state traversal is 98 % writes)
    st_file = os.path.join(ROOT, "profiles", "r1s_hbm_streams.json")
    tb = line["roofline"].get("traffic")
    if tb and partial_ms and partial_launches and os.path.exists(st_file) and n == 4:
        try:
            st = json.load(open(st_file))
            gbs = tb / (partial_ms / partial_launches * 1e-3) / 1e9
            line["roofline"]["physical"] = {"dram_gbs": gbs, "write_peak_gbs": st["write_t32_gbs"],
                                            "frac_of_write_peak": gbs / st["write_t32_gbs"],
                                            "peak_source": "profiles/r1s_hbm_streams.json (write-only stream in the kernel's store shape)"}
        except Exception:
            pass
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
