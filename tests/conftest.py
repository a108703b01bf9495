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


DAEMON_BIN = os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200", "bin", "gpushare-device-plugin-v2")
INSPECT_BIN = os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200", "bin", "kubectl-inspect-gpushare-v2")
CSRC = os.path.join(ROOT, "aliyun-gpushare-device-plugin_b200", "csrc")
HOST_SOURCES = ("allocate.cpp", "table.cpp", "gosem.cpp", "plugin.cpp", "errors.cpp", "grpc_server.cpp",
                "grpc_client.cpp")
_SANITIZED = {}


def sanitized_objects(kind):
    """The host sources of the core library compiled once per session with a sanitizer
    ("asan": address+undefined, "tsan": thread), in parallel; the sanitizer tests link their
    drivers against these objects instead of each recompiling the library."""
    if kind in _SANITIZED:
        return _SANITIZED[kind]
    import tempfile
    from concurrent.futures import ThreadPoolExecutor
    flags = {"asan": ["-fsanitize=address,undefined", "-fno-sanitize-recover=all"],
             "tsan": ["-fsanitize=thread"]}[kind]
    out_dir = tempfile.mkdtemp(prefix="gs_%s_" % kind)

    def compile_one(name):
        obj = os.path.join(out_dir, name.replace(".cpp", ".o"))
        subprocess.run(["g++", "-std=c++17", "-O1", "-g"] + flags + ["-I" + os.path.join(ROOT, "include"),
                       "-I" + CSRC, "-c", os.path.join(CSRC, name), "-o", obj], check=True)
        return obj
    with ThreadPoolExecutor(max_workers=len(HOST_SOURCES)) as pool:
        objs = list(pool.map(compile_one, HOST_SOURCES))
    _SANITIZED[kind] = (flags, objs)
    return _SANITIZED[kind]


def build_sanitized(kind, driver, exe):
    flags, objs = sanitized_objects(kind)
    subprocess.run(["g++", "-std=c++17", "-O1", "-g"] + flags + ["-I" + os.path.join(ROOT, "include"), "-I" + CSRC,
                   driver] + objs + ["-o", str(exe), "-lpthread"], check=True)
    return str(exe)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # Build once if the shared objects are missing (fresh checkout); never on the GPU box
    # if they travelled with the snapshot.
    if not all(os.path.exists(p) for p in (FAKE_NVML, ORACLE_SO, CORE_SO, DAEMON_BIN, INSPECT_BIN)):
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
def pattern_kat_hash():
    with open(os.path.join(ROOT, "tests", "golden", "pattern_kat.json")) as f:
        return json.load(f)["hash_vectors"]


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
