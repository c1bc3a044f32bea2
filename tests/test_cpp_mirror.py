"""The C++ host mirror (include/b200crc.hpp): compiles everywhere, runs against the oracle on a GPU."""
import importlib
import os
import subprocess

import pytest

import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
EXE = os.path.join(ROOT, "tests", "cpp", "host_mirror_test")


def build_exe():
    importlib.import_module("custom-go-client-benchmark_b200.build").build()
    oracle.build()
    libdir = os.path.dirname(pkg._lib.SO_PATH)
    odir = os.path.join(ROOT, "oracle")
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-I", os.path.join(ROOT, "include"), "-I", odir,
           os.path.join(ROOT, "tests", "cpp", "host_mirror_test.cpp"), "-o", EXE,
           "-L", libdir, "-lb200crc", "-L", odir, "-loracle", f"-Wl,-rpath,{libdir}", f"-Wl,-rpath,{odir}"]
    subprocess.check_call(cmd)
    return EXE


def test_cpp_mirror_compiles_and_links():
    assert os.path.exists(build_exe())


@pytest.mark.gpu
def test_cpp_mirror_matches_oracle():
    exe = build_exe()
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr + out.stdout
    assert "host mirror ok" in out.stdout
