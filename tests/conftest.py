import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FAKE_NVML = os.path.join(ROOT, "tests", "fake_nvml", "libfake_nvml.so")
ORACLE_SO = os.path.join(ROOT, "oracle", "libgs_oracle.so")
CORE_SO = os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200", "libgpushare_core.so")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # Build once if the shared objects are missing (fresh checkout); never on the GPU box
    # if they travelled with the snapshot.
    if not all(os.path.exists(p) for p in (FAKE_NVML, ORACLE_SO, CORE_SO)):
        subprocess.run(["make", "-C", ROOT, "-j4"], check=True, stdout=subprocess.DEVNULL)


def pytest_collection_modifyitems(config, items):
    # keep tests/test_zz_* last whatever order files were given in
    items.sort(key=lambda it: os.path.basename(str(it.fspath)).startswith("test_zz_"))


@pytest.fixture(scope="session")
def golden_cases():
    with open(os.path.join(ROOT, "tests", "golden", "allocate_cases.json")) as f:
        return json.load(f)["cases"]


@pytest.fixture(scope="session")
def pattern_kat():
    with open(os.path.join(ROOT, "tests", "golden", "pattern_kat.json")) as f:
        return json.load(f)["vectors"]


@pytest.fixture(scope="session")
def oracle_lib():
    import ctypes
    return ctypes.CDLL(ORACLE_SO)


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False
