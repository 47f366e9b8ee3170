"""CPU: the C-ABI library loads, exports every symbol include/fcp.h declares,
and its host-only arithmetic (finishing raw CRC values) is right.  No kernel
is launched here."""
import os
import re

import numpy as np

from dlrover_wuqiong_b200 import _native
from oracle import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    inc = os.path.join(ROOT, "include")
    for f in os.listdir(inc):
        if f.endswith(".h"):
            txt = open(os.path.join(inc, f)).read()
            txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
            names |= set(re.findall(r"\b(fcp_[a-z0-9_]+)\s*\(", txt))
    return names


def test_library_exports_every_declared_symbol():
    lib = _native.lib()
    declared = _declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), name
    bound = {s[0] for s in _native.SYMBOLS}
    assert declared == bound


def test_version_and_error_string():
    lib = _native.lib()
    assert lib.fcp_version() == 1
    assert isinstance(lib.fcp_last_error(), bytes)


def _raw_crc(data):
    # init 0, no final xor, bytewise
    c = 0
    for b in data:
        c ^= b
        for _ in range(8):
            c = (c >> 1) ^ (0x82F63B78 if c & 1 else 0)
    return c


def test_finish_checksums_strips_padding_and_applies_init():
    rng = np.random.default_rng(11)
    chunk = 16384
    for total in (1, 15, 16, 100, 16384, 16385, 2 * 16384 + 777):
        buf = rng.integers(0, 256, total, dtype=np.uint8)
        n = (total + chunk - 1) // chunk
        raw = []
        for i in range(n):
            part = buf[i * chunk:(i + 1) * chunk].tobytes()
            part += bytes(-len(part) % _native.REGION_BYTES)
            raw.append(_raw_crc(part))
        got = _native.finish_checksums(np.array(raw, dtype=np.uint32), total, chunk)
        want = oracle.chunk_crc32c(buf, chunk)
        assert got.tolist() == want.tolist(), total


def test_bad_geometry_is_rejected():
    import pytest
    with pytest.raises(_native.FcpError):
        _native.finish_checksums(np.zeros(1, dtype=np.uint32), 10, 1000)
