"""CPU: the C-ABI library loads and exports every symbol include/mcpt.h declares; compute entry points fail loudly
without a GPU (there is no CPU fallback)."""
import ctypes
import os
import re

import pytest

import montecarlopathtracing_b200 as M
from montecarlopathtracing_b200 import _capi
from conftest import ROOT


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "mcpt.h")).read()
    return sorted(set(re.findall(r"MCPT_API\s+[\w\s\*]+?\b(mcpt_\w+)\s*\(", hdr)))


def test_header_symbols_are_exported_and_bound(built_lib):
    names = _declared_symbols()
    assert len(names) >= 20
    lib = ctypes.CDLL(built_lib)
    for n in names:
        assert hasattr(lib, n), "libmcpt.so does not export " + n
    assert sorted(_capi.SIGNATURES) == names, "python binding and header disagree"


def test_struct_sizes_match_header(built_lib):
    assert ctypes.sizeof(_capi.Material) == 72 and ctypes.sizeof(_capi.Light) == 32 and ctypes.sizeof(_capi.Camera) == 88
    assert ctypes.sizeof(_capi.Hit) == 20 and ctypes.sizeof(_capi.Stats) == 64 and ctypes.sizeof(_capi.SceneDesc) == 72


def test_no_cpu_fallback(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the loud-failure path is for GPU-less hosts")
    with pytest.raises(M.McptError, match="no CPU fallback"):
        M.Context(0)
    with pytest.raises(M.McptError):
        M.render_scene(os.path.join(ROOT, "scenes") + "/", "cornell-box", 1, out_dir="/tmp/")


def test_product_never_imports_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may touch oracle/."""
    pkg = os.path.join(ROOT, "montecarlopathtracing_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath.split(os.sep)[-1:]:
            continue
        for f in files:
            if f.endswith((".py", ".cpp", ".h", ".cu", ".cuh")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r'#include\s*[<"][^>"]*oracle|import\s+oracle|from\s+oracle|oracle_py|liboracle', text), f
                # the only run-time loading the product does is NCCL for the multi-GPU exchange (multigpu.cu), by literal name
                for call in re.findall(r'dlopen\s*\(([^;]*);', text):
                    assert '"libnccl.so' in call, (f, call)
