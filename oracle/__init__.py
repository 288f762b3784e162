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
