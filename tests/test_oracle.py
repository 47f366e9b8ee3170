"""CPU: the oracle against the reference's golden vectors and the published
known-answer tests of the checksum algorithms."""
import numpy as np
import pytest
import torch

from oracle import oracle


def _meta_equal(ref_meta, o_meta):
    """ref_meta: plain dict from the fixture; o_meta: oracle tree."""
    if isinstance(ref_meta, dict) and ref_meta.get("__tensor_meta__"):
        assert isinstance(o_meta, oracle.OMeta)
        assert tuple(ref_meta["shape"]) == tuple(o_meta.shape)
        assert ref_meta["dtype"] == o_meta.dtype
        assert ref_meta["element_size"] == o_meta.element_size
        assert ref_meta["numel"] == o_meta.numel
        assert ref_meta["offset"] == o_meta.offset
        return
    if isinstance(ref_meta, dict):
        keys = [k for k in ref_meta if k != "_DLORVER_CKPT_CONFIG"]
        assert keys == list(o_meta.keys())
        for k in keys:
            _meta_equal(ref_meta[k], o_meta[k])
    elif isinstance(ref_meta, list):
        assert len(ref_meta) == len(o_meta)
        for a, b in zip(ref_meta, o_meta):
            _meta_equal(a, b)
    elif isinstance(ref_meta, tuple):
        assert isinstance(o_meta, tuple) and len(ref_meta) == len(o_meta)
    else:
        assert ref_meta == o_meta


def test_crc32c_kat():
    # RFC 3720 B.4 vectors
    assert oracle.crc32c(b"123456789") == 0xE3069283
    assert oracle.crc32c(bytes(32)) == 0x8A9136AA
    assert oracle.crc32c(b"\xff" * 32) == 0x62A8AB43
    assert oracle.crc32c(bytes(range(32))) == 0x46DD794E
    assert oracle.crc32c(b"") == 0


def test_xxh64_kat():
    assert oracle.xxh64(b"") == 0xEF46DB3751D8E999
    xxhash = pytest.importorskip("xxhash")
    rng = np.random.default_rng(7)
    for n in [1, 3, 4, 7, 8, 31, 32, 33, 63, 64, 1000, 4097]:
        data = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        for seed in (0, 1, 2**63 + 5):
            assert oracle.xxh64(data, seed) == xxhash.xxh64(data, seed=seed).intdigest()


def test_golden_sizes(golden):
    assert golden["cases"]["simplenet_sgd_9640"]["size"] == 9640
    assert golden["cases"]["toymodel_adam_1632"]["size"] == 1632


@pytest.mark.parametrize("name", ["simplenet_sgd_9640", "toymodel_adam_1632", "quirks",
                                  "mlp_adamw_stepped", "noncontig"])
def test_oracle_matches_reference(golden, name):
    case = golden["cases"][name]
    meta, total, buf = oracle.pack(case["state_dict"])
    assert total == case["size"]
    _meta_equal(case["meta"], meta)
    assert buf.tobytes() == case["shm"]
    # restore half: views equal inputs
    back = oracle.unpack(meta, buf)
    flat_in, _, _ = oracle.flatten(case["state_dict"])
    flat_out, _, _ = oracle.flatten(back)
    assert len(flat_in) == len(flat_out)
    for (a, _), (b, _) in zip(flat_in, flat_out):
        assert a.dtype == b.dtype
        if a.numel() == 0:  # SAVER:124-125: empty tensors come back as shape (0,)
            assert tuple(b.shape) == (0,)
            continue
        assert tuple(a.shape) == tuple(b.shape)
        assert oracle.tensor_bytes(a).tobytes() == oracle.tensor_bytes(b).tobytes()


def test_quirk_offsets(golden):
    # dense offsets 0/48/52/55/71/87/90 ...
    flat, _, _ = oracle.flatten(golden["cases"]["quirks"]["state_dict"])
    assert [m.offset for _, m in flat][:7] == [0, 48, 52, 55, 71, 87, 90]


def test_second_save_reuses_offsets(golden):
    case = golden["cases"]["mlp_adamw_stepped"]
    meta, total, buf = oracle.pack(case["second"]["state_dict"])
    assert buf.tobytes() == case["second"]["shm"]
    assert case["second"]["meta"]["step"] == 2


def test_c_pack_equals_python_pack(golden):
    sd = golden["cases"]["mlp_adamw_stepped"]["state_dict"]
    flat, _, total = oracle.flatten(sd)
    keep = [oracle.tensor_bytes(t).copy() for t, _ in flat]
    d = np.zeros(len(flat), dtype=oracle.DESC_DTYPE)
    for i, ((_, m), k) in enumerate(zip(flat, keep)):
        d[i] = (k.ctypes.data if k.size else 0, m.offset, k.size)
    out = oracle.pack_descs(d[d["nbytes"] > 0], total)
    assert out.tobytes() == golden["cases"]["mlp_adamw_stepped"]["shm"]


def test_chunk_crc():
    rng = np.random.default_rng(3)
    buf = rng.integers(0, 256, 5 * 4096 + 123, dtype=np.uint8)
    sums = oracle.chunk_crc32c(buf, 4096)
    assert len(sums) == 6
    assert int(sums[5]) == oracle.crc32c(buf[5 * 4096:])
    assert int(sums[0]) == oracle.crc32c(buf[:4096])
