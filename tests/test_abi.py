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


def test_save_leg_tuner_state_machine():
    """The run-time choice of the device->host leg (transfer.SaveLegTuner): no
    GPU needed, it only looks at (mode, bytes, seconds)."""
    from dlrover_wuqiong_b200 import _native, transfer

    G = 1 << 30
    CE, K = _native.MODE_STAGED, _native.MODE_STAGED_KERNEL
    # healthy link: the first save is ignored, the second settles on the copy engine
    t = transfer.SaveLegTuner()
    assert t.mode() == CE
    t.report(CE, 19 * G, 0.9)  # first save of a process: slow, not a sample
    assert not t.settled and t.mode() == CE
    t.report(CE, 19 * G, 19 * G / 57e9)
    assert t.settled and t.choice == CE and t.mode() == CE
    # copy-engine writes do not scale on this host: try the kernel, keep it if > 10 % faster
    t = transfer.SaveLegTuner()
    t.report(CE, 19 * G, 1.0)
    t.report(CE, 19 * G, 19 * G / 33e9)
    assert not t.settled and t.mode() == K
    t.report(K, 19 * G, 19 * G / 51e9)
    assert t.settled and t.choice == K and t.mode() == K
    # ... or not
    t = transfer.SaveLegTuner()
    t.report(CE, 19 * G, 1.0)
    t.report(CE, 19 * G, 19 * G / 33e9)
    t.report(K, 19 * G, 19 * G / 34e9)
    assert t.settled and t.choice == CE
    # small saves say nothing about the link; a failing kernel leg ends the probe
    t = transfer.SaveLegTuner()
    for _ in range(5):
        t.report(CE, 100 << 20, 1.0)
    assert t.state == "warmup"
    t.report(CE, 2 * G, 1.0)
    t.report(CE, 2 * G, 1.0)
    assert t.mode() == K
    t.kernel_leg_failed()
    assert t.settled and t.mode() == CE
