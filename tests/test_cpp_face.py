"""include/gpushare.hpp compiles and behaves: a C++ consumer runs golden case G1."""
import os
import subprocess

from tests.conftest import CORE_SO, ROOT, build_sanitized


def test_cpp_consumer(tmp_path):
    exe = tmp_path / "test_plugin"
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-I" + os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "cpp", "test_plugin.cpp"), CORE_SO, "-o", str(exe),
                    "-Wl,-rpath," + os.path.dirname(CORE_SO)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0 and "CPP FACE OK" in out.stdout, out.stdout + out.stderr


def test_host_core_under_address_and_ub_sanitizers(tmp_path):
    """allocate/table/gosem/plugin/errors rebuilt with ASan + UBSan and driven with random
    tables, health marks, pods and (often malformed) requests: no out-of-bounds, overflow or leak."""
    exe = build_sanitized("asan", os.path.join(ROOT, "tests", "cpp", "fuzz_core.cpp"), tmp_path / "fuzz_core")
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "FUZZ CORE OK" in out.stdout, (out.stdout + out.stderr)[-3000:]


def test_host_core_under_thread_sanitizer(tmp_path):
    """Allocate from four threads, health marks, two ListAndWatch senders and the health wait, all
    on one table/plugin, under TSan: no data race."""
    exe = build_sanitized("tsan", os.path.join(ROOT, "tests", "cpp", "race_core.cpp"), tmp_path / "race_core")
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "RACE CORE OK" in out.stdout and "WARNING: ThreadSanitizer" not in out.stderr, \
        (out.stdout + out.stderr)[-4000:]
