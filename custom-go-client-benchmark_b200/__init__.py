"""B200-native CRC-32C verification and latency reduction: the data-parallel hot path of
tritone/custom-go-client-benchmark behind a C ABI (include/b200crc.h).

    crc32c   — Go hash/crc32-shaped surface (New/Update/Checksum/VerifyObjects/Combine)
    latency  — README histogram, OpenCensus distribution, ssd_test percentiles
    readloop — the reference's ReadObject loop (main.go:119-156) over an in-memory reader (BASELINE config 1)
    blockread — the reference's local-SSD block-read benchmark (benchmark-script/ssd_test/main.go)
The CUDA library is required; nothing here computes on the CPU.
"""
from . import _lib, blockread, crc32c, latency, readloop  # noqa: F401
from ._lib import B200Error  # noqa: F401
