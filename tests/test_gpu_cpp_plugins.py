"""The C++ host layer (include/fgs_plugins.hpp: the reference's plugin interface over the C ABI) driven by the
reference's own test cases — tests/cpp/plugin_test.cpp. The Go toolchain is absent on this image; this is the
host side a maintainer would write in Go over cgo, in C++."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "flex-gpu-scheduler_b200")
OUT = os.path.join(ROOT, "tests", "cpp", "_build")


def _build():
    os.makedirs(OUT, exist_ok=True)
    exe = os.path.join(OUT, "plugin_test")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(PKG, "csrc"),
           "-I", os.path.join(ROOT, "tests", "cpp"), os.path.join(ROOT, "tests", "cpp", "plugin_test.cpp"), "-o", exe,
           "-L", PKG, "-lfgs", "-Wl,-rpath," + PKG]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_cpp_host_layer_compiles_and_links():
    # CPU check: the header-only plugin layer and the fixtures compile against include/fgs.h and link with libfgs.so
    assert os.path.exists(_build())


@pytest.mark.gpu
def test_reference_plugin_tests_through_the_cpp_host_layer(gpu_ok):
    exe = _build()
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failed" in r.stdout
