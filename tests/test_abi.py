"""The C-ABI library loads without a GPU and exports every symbol include/gpushare.h declares."""
import ctypes
import os
import re
import subprocess

from gpushare_b200 import _lib
from tests.conftest import CORE_SO, ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "gpushare.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"GS_API\s+[\w\s\*]+?\b(gs_\w+)\s*\(", text)))


def test_header_symbols_all_exported_and_bound():
    names = declared_symbols()
    assert len(names) >= 60
    lib = ctypes.CDLL(CORE_SO)
    for n in names:
        assert hasattr(lib, n), "not exported: " + n
    assert sorted(_lib.SIGNATURES) == names          # the ctypes table mirrors the header
    _lib.load()


def test_no_unexpected_exports_and_no_oracle_linkage():
    out = subprocess.run(["nm", "-D", "--defined-only", CORE_SO], capture_output=True, text=True,
                         check=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert {e for e in exported if e.startswith("gs_")} == set(declared_symbols())
    assert not [e for e in exported if "oracle" in e or e.startswith("ref_")]
    ldd = subprocess.run(["ldd", CORE_SO], capture_output=True, text=True).stdout
    assert "gs_oracle" not in ldd and "libcuda.so" not in ldd and "nvidia-ml" not in ldd


def test_product_sources_never_touch_oracle():
    pkg = os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "ref_model" not in src and "libgs_oracle" not in src, f
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f


def test_status_strings_and_version():
    lib = _lib.load()
    assert lib.gs_strerror(0) == b"ok" and b"buffer" in lib.gs_strerror(_lib.GS_E_NOSPACE)
    assert b"sm_100a" in lib.gs_version()


def test_probe_without_gpu_fails_loudly():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("has a GPU")
    from gpushare_b200.probe import Probe
    with pytest.raises(_lib.GpushareError) as e:
        Probe(0)
    assert e.value.status == _lib.GS_E_CUDA
