"""include/gpushare.hpp compiles and behaves: a C++ consumer runs golden case G1."""
import os
import subprocess

from tests.conftest import CORE_SO, ROOT


def test_cpp_consumer(tmp_path):
    exe = tmp_path / "test_plugin"
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-I" + os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "cpp", "test_plugin.cpp"), CORE_SO, "-o", str(exe),
                    "-Wl,-rpath," + os.path.dirname(CORE_SO)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0 and "CPP FACE OK" in out.stdout, out.stdout + out.stderr
