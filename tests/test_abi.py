"""The C-ABI library loads and exports every symbol include/b200lk.h declares (no GPU needed)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200lk.h")).read()
    return sorted(set(re.findall(r"\b(b200lk_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_header_symbols():
    from iqtree2_b200 import build
    lib = ctypes.CDLL(build.build())
    names = declared_symbols()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), name


def test_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from iqtree2_b200.capi import B200LKError, Engine
    with pytest.raises(B200LKError):
        Engine(ntaxa=4, nstates=4, ncat=4, nptn=10, nslots=4, n_state_codes=19)


def test_oracle_mirrors_abi(oracle_lib):
    lib = ctypes.CDLL(oracle_lib)
    for name in ("create", "destroy", "upload_tips", "set_ptn_invar", "set_model", "update_partials",
                 "branch_lnl", "build_theta", "derv", "lnl_from_buffer", "evaluate"):
        assert hasattr(lib, "orc_" + name)
