"""include/b200crc.h as PLAIN C (what cgo compiles it as): tests/c/abi_test.c builds with gcc -std=c11 -pedantic
everywhere and, on a GPU, passes its known answers / Hash32 / VerifyObjects checks through the C ABI."""
import importlib
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pkg = importlib.import_module("custom-go-client-benchmark_b200")
EXE = os.path.join(ROOT, "tests", "c", "abi_test")


def build_exe():
    importlib.import_module("custom-go-client-benchmark_b200.build").build()
    libdir = os.path.dirname(pkg._lib.SO_PATH)
    cmd = ["gcc", "-std=c11", "-O1", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "c", "abi_test.c"), "-o", EXE, "-L", libdir, "-lb200crc", f"-Wl,-rpath,{libdir}"]
    subprocess.check_call(cmd)
    return EXE


def test_c_caller_compiles_and_links():
    assert os.path.exists(build_exe())


@pytest.mark.gpu
def test_c_caller_known_answers_and_verify_objects():
    exe = build_exe()
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr + out.stdout
    assert "abi_test ok" in out.stdout
