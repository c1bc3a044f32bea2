"""CPU-side checks of the drop-in boundary: the library loads, exports every symbol include/b200crc.h
declares, and fails loudly (no CPU fallback) when there is no GPU."""
import ctypes as C
import importlib
import os
import re
import subprocess

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pkg = importlib.import_module("custom-go-client-benchmark_b200")


@pytest.fixture(scope="module")
def lib():
    importlib.import_module("custom-go-client-benchmark_b200.build").build()
    return pkg._lib.load()


def declared_functions():
    text = open(os.path.join(ROOT, "include", "b200crc.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    names = declared_functions()
    for must in ("b200_init", "b200_crc32c_update_host", "b200_hasher_write", "b200_crc32c_batch_device",
                 "b200_crc32c_batch_host", "b200_crc32c_sharded", "b200_crc32c_stream_host", "b200_crc32c_combine",
                 "b200_hist_fixed", "b200_select_kth"):
        assert must in names


def test_library_exports_every_declared_symbol(lib):
    out = subprocess.check_output(["nm", "-D", "--defined-only", pkg._lib.SO_PATH], text=True)
    exported = set(re.findall(r"\sT\s+(b200_\w+)", out))
    for name in declared_functions():
        assert name in exported, f"{name} declared in include/b200crc.h but not exported"
        assert hasattr(lib, name)


def test_ctypes_table_matches_header():
    assert sorted(pkg._lib.SIGNATURES) == declared_functions()


def test_no_torch_or_cxx_types_in_signatures():
    text = open(os.path.join(ROOT, "include", "b200crc.h")).read()
    assert "torch" not in text and "std::" not in text and "at::" not in text


def test_product_does_not_link_or_import_the_oracle():
    out = subprocess.check_output(["ldd", pkg._lib.SO_PATH], text=True)
    assert "oracle" not in out
    for dirpath, _, files in os.walk(os.path.join(ROOT, "custom-go-client-benchmark_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".go")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in src and "liboracle" not in src and "oracle_" not in src, f


def test_strerror(lib):
    assert lib.b200_strerror(0) == b"ok"
    assert lib.b200_strerror(-1) == b"invalid argument"
    assert lib.b200_strerror(-4) == b"empty sample set"


# b200_crc32c_combine is pure host GF(2) arithmetic (no message bytes): it may run without a GPU.
@settings(max_examples=50, deadline=None)
@given(st.binary(max_size=2000), st.binary(max_size=2000))
def test_combine_matches_oracle(a, b):
    lib = pkg._lib.load()
    c = lib.b200_crc32c_combine(oracle.crc32c(a), oracle.crc32c(b), len(b))
    assert c == oracle.crc32c(a + b)


def test_combine_large_shifts():
    for n in (0, 1, 511, 512, 1 << 21, (1 << 33) + 5, 1 << 36, (1 << 40) - 1):
        assert pkg.crc32c.Combine(0xDEADBEEF, 0x12345678, n) == oracle.combine(0xDEADBEEF, 0x12345678, n)


def _has_gpu(lib):
    n = C.c_int()
    return lib.b200_device_count(C.byref(n)) == 0 and n.value > 0


def test_fails_loudly_without_gpu(lib):
    if _has_gpu(lib):
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = lib.b200_init(1, None, C.byref(h))
    assert rc == pkg._lib.ERR_CUDA
    assert lib.b200_last_error()
    with pytest.raises(pkg.B200Error):
        pkg.crc32c.Checksum(b"123456789")


def test_missing_library_is_an_error(monkeypatch):
    monkeypatch.setattr(pkg._lib, "_lib", None)
    monkeypatch.setattr(pkg._lib, "SO_PATH", "/nonexistent/libb200crc.so")
    with pytest.raises(FileNotFoundError):
        pkg._lib.load()


def test_wire_checksum_codecs():
    import base64
    c = pkg.crc32c
    crc = 0xE3069283  # CRC-32C("123456789")
    hdr = c.EncodeXGoogHash(crc)
    assert hdr == "crc32c=" + base64.b64encode(bytes([0xE3, 0x06, 0x92, 0x83])).decode() == "crc32c=4waSgw=="
    assert c.DecodeXGoogHash(hdr) == crc
    assert c.DecodeXGoogHash("md5=XrY7u+Ae7tCTyyK7j1rNww==, crc32c=4waSgw==") == crc
    assert c.DecodeXGoogHash("md5=XrY7u+Ae7tCTyyK7j1rNww==") is None
    with pytest.raises(ValueError):
        c.DecodeXGoogHash("crc32c=AAAA")
    assert c.DecodeFixed32(c.EncodeFixed32(crc)) == crc and c.EncodeFixed32(crc) == bytes([0x83, 0x92, 0x06, 0xE3])
    assert oracle.crc32c(b"123456789") == crc


def test_distribution_export_shape():
    d = pkg.latency.Distribution(count=4, min=1, max=7, sum=16, sum_wrapped=16, sumsq=84,
                                 counts=np.array([1, 2, 1], dtype=np.uint64))
    oc = d.to_opencensus(bounds=[2, 5])
    assert oc["Count"] == 4 and oc["Mean"] == 4.0 and oc["SumOfSquaredDev"] == 84 - 16 * 16 / 4
    assert oc["CountPerBucket"] == [1, 2, 1] and oc["Min"] == 1.0 and oc["Max"] == 7.0
    assert d.avg_go_int == 4
    neg = pkg.latency.Distribution(3, -5, 1, -7, -7, 27, np.zeros(1, dtype=np.uint64))
    assert neg.avg_go_int == -2  # Go: -7 / 3 == -2 (truncation toward zero)
