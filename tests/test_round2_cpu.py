"""CPU: host-side robustness added in round 2 — the shm populator never writes
into a segment, writing_shm=True is published before the first byte moves, the
agent re-checks the recorded chunk CRCs before it persists a shard, raw image
files carry the sums of their own payload, and the host CRC-32C routine of
libfcp.so (pure host code) agrees with the oracle."""
import os
import tempfile

import numpy as np
import pytest
import torch

os.environ.setdefault("ROLE_NAME", "dlrover-trainer")
os.environ.setdefault("TORCHELASTIC_RUN_ID", "fcp_test_%d" % os.getpid())

from dlrover_wuqiong_b200 import _native, raw_ckpt  # noqa: E402
from dlrover_wuqiong_b200.ckpt_saver import (  # noqa: E402
    DLROVER_CKPT_CONFIG_KEY,
    FCP_CHECKSUM_KEY,
    CheckpointConfig,
    CommonDirCheckpointSaver,
    SharedMemoryHandler,
)
from dlrover_wuqiong_b200.common import multi_process  # noqa: E402
from dlrover_wuqiong_b200.common.multi_process import SharedMemory  # noqa: E402
from dlrover_wuqiong_b200.common.serialize import ClassMeta  # noqa: E402
from dlrover_wuqiong_b200.common.storage import PosixDiskStorage  # noqa: E402
from oracle import oracle  # noqa: E402


# ------------------------------------------------------------ host CRC-32C
@pytest.mark.parametrize("n", [0, 1, 7, 8, 9, 4095, 3 * 4096, 3 * 4096 + 5, (3 << 20) + 123])
def test_host_crc32c_matches_oracle(n):
    rng = np.random.default_rng(n)
    a = rng.integers(0, 256, n + 3, dtype=np.uint8)
    for off in (0, 1, 3):  # unaligned starts
        b = a[off:off + n]
        want = oracle.chunk_crc32c(b, 65536)
        got = _native.host_chunk_crc32c(b.ctypes.data, b.size, 65536, threads=3) \
            if n else np.zeros(0, np.uint32)
        assert got.tolist() == want.tolist()


def test_host_crc32c_known_answer():
    # RFC 3720 B.4
    a = np.frombuffer(b"123456789", dtype=np.uint8)
    assert _native.host_chunk_crc32c(a.ctypes.data, 9, 1 << 20).tolist() == [0xE3069283]
    z = np.zeros(32, dtype=np.uint8)
    assert _native.host_chunk_crc32c(z.ctypes.data, 32, 1 << 20).tolist() == [0x8A9136AA]


def test_oracle_hw_crc_matches_bytewise():
    rng = np.random.default_rng(7)
    a = rng.integers(0, 256, (5 << 20) + 77, dtype=np.uint8)
    assert oracle.chunk_crc32c_fast(a, 1 << 20).tolist() == oracle.chunk_crc32c(a, 1 << 20).tolist()
    assert oracle.chunk_crc32c_fast(a[:5], 1 << 20, threads=4).tolist() == \
        [oracle.crc32c(a[:5])]


# --------------------------------------------------------------- populator
def test_populator_fallback_never_writes(monkeypatch):
    """On a kernel without MADV_POPULATE_WRITE the populate threads must leave
    the pages alone: the owner may already be filling the segment."""
    monkeypatch.setattr(multi_process, "_MADV_POPULATE_WRITE", 9999)  # EINVAL
    monkeypatch.setattr(multi_process, "_POPULATE_BLOCK", 1 << 20)
    name = "fcp_t_pop_%d" % os.getpid()
    size = 48 << 20
    shm = SharedMemory(name=name, create=True, size=size)
    try:
        view = np.frombuffer(shm.buf, dtype=np.uint8)
        view[:] = 0xA5  # races with the populate threads on purpose
        shm.wait_populated()
        assert int((view != 0xA5).sum()) == 0
        del view
    finally:
        shm.close()
        shm.unlink()


def test_host_save_with_forced_populate_fallback(monkeypatch):
    monkeypatch.setattr(multi_process, "_MADV_POPULATE_WRITE", 9999)
    monkeypatch.setattr(multi_process, "_POPULATE_BLOCK", 1 << 20)
    h = SharedMemoryHandler(37, host=True)
    try:
        sd = {"w": [torch.full((1 << 20,), float(i)) for i in range(12)],
              DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=3, paths={})}
        h.save_state_dict(sd)
        back = h.load_state_dict()
        for i, t in enumerate(back["w"]):
            assert torch.equal(t, sd["w"][i])
    finally:
        h.close()
        h.unlink()


# -------------------------------------------------- writing_shm before bytes
def test_writing_shm_is_published_before_the_copy(monkeypatch):
    from dlrover_wuqiong_b200 import ckpt_saver
    h = SharedMemoryHandler(38, host=True)
    seen = []
    real = ckpt_saver._host_copy_leaves

    def spy(leaves, buf):
        # what a reader of the agent's meta would see at this moment
        seen.append(h.metadata.get()[DLROVER_CKPT_CONFIG_KEY].writing_shm)
        return real(leaves, buf)

    monkeypatch.setattr(ckpt_saver, "_host_copy_leaves", spy)
    try:
        for step in (1, 2):
            h.save_state_dict({"w": torch.ones(100),
                               DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=step, paths={})})
        assert seen == [True, True]
        assert h.metadata.get()[DLROVER_CKPT_CONFIG_KEY].writing_shm is False
    finally:
        h.close()
        h.unlink()


# ------------------------------------------------------ agent-side verify
class _Saver(CommonDirCheckpointSaver):
    pass


def _fake_gpu_image(handler, step, path):
    """A host-written image that carries chunk sums like a GPU save would."""
    sd = {"model_states": {"a": torch.arange(300000, dtype=torch.float32),
                           "b": torch.arange(7, dtype=torch.int64)},
          DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=step, paths={"model_states": path})}
    handler.save_state_dict(sd)
    meta = handler.metadata.get()
    total = handler.shared_memory.size
    sums = oracle.chunk_crc32c(np.frombuffer(handler.shared_memory.buf, dtype=np.uint8), 65536)
    meta[FCP_CHECKSUM_KEY] = {"algo": "crc32c", "chunk_bytes": 65536, "sums": sums.tolist()}
    handler.metadata.set(meta)
    return sd, total


def test_agent_refuses_to_persist_a_corrupt_image():
    with tempfile.TemporaryDirectory() as tmp:
        saver = _Saver(tmp, ClassMeta(PosixDiskStorage.__module__, "PosixDiskStorage", {}),
                       local_shard_num=1, global_shard_num=1)
        try:
            h = saver._shm_handlers[0]
            path = os.path.join(tmp, "10", "rank_0.pt")
            sd, total = _fake_gpu_image(h, 10, path)
            assert h.verify_image() == (total + 65535) // 65536
            conf = h.get_checkpoint_config(CheckpointConfig())
            os.makedirs(os.path.dirname(path))
            saver.persist_to_storage(0, conf)  # intact: written
            assert torch.equal(torch.load(path)["a"], sd["model_states"]["a"])
            os.unlink(path)
            h.shared_memory.buf[70000] ^= 0x40  # one flipped bit in chunk 1
            with pytest.raises(_native.ChecksumMismatch, match="chunk 1"):
                saver.persist_to_storage(0, conf)
            assert not os.path.exists(path)
            # _save_shard reports the failure instead of dropping a done-file
            done = os.path.join(tmp, "done")
            os.makedirs(done)
            assert saver._save_shard(10, 0, conf, done) is False
            assert os.listdir(done) == []
        finally:
            saver.close()


# -------------------------------------------------- raw files: payload sums
def test_raw_file_records_sums_of_its_own_payload(monkeypatch):
    monkeypatch.setenv(raw_ckpt.PERSIST_FORMAT_ENV, "raw")
    h = SharedMemoryHandler(39, host=True)
    with tempfile.TemporaryDirectory() as tmp:
        try:
            # two state names: the second one's byte range starts mid-chunk
            sd = {"model_states": {"a": torch.arange(50000, dtype=torch.float32)},
                  "optim_states": {"m": torch.arange(90001, dtype=torch.float32) * 0.5,
                                   "step": 7},
                  DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=4, paths={})}
            h.save_state_dict(sd)
            meta = h.metadata.get()
            image = np.frombuffer(h.shared_memory.buf, dtype=np.uint8)
            meta[FCP_CHECKSUM_KEY] = {"algo": "crc32c", "chunk_bytes": 65536,
                                      "sums": oracle.chunk_crc32c(image, 65536).tolist()}
            h.metadata.set(meta)
            for name, img in h.load_raw_images().items():
                raw_ckpt.write_raw(img, os.path.join(tmp, name))
            for name in ("model_states", "optim_states"):
                header, data_off = raw_ckpt.read_header(os.path.join(tmp, name))
                lo, hi = raw_ckpt.subtree_range(meta[name])
                want = oracle.chunk_crc32c(image[lo:hi], 65536).tolist()
                assert header["payload_checksums"]["sums"] == want
                assert header["image"]["total"] == image.size
            back = raw_ckpt.read_raw(os.path.join(tmp, "optim_states"))
            assert torch.equal(back["m"], sd["optim_states"]["m"]) and back["step"] == 7
            del image
        finally:
            h.close()
            h.unlink()


# ------------------------------------------------ peer restore: who provides
def test_plan_providers():
    from dlrover_wuqiong_b200.flash_checkpoint.peer_restore import plan_providers
    MB = 1 << 20
    total, chunk = 10 * MB + 5, 2 * MB  # 6 chunks

    def f(host, hbm=(0, 0, 0), shm=0):
        return {"host": host, "hbm": hbm, "shm_step": shm}

    # restart: nobody holds HBM, every rank stages its stripe from shm
    got = plan_providers([f("a", shm=5), f("a", shm=5), f("a", shm=5)], 5, total, chunk)
    assert got == {"a": [(0, 4 * MB, 0, True), (4 * MB, 8 * MB, 1, True),
                         (8 * MB, total, 2, True)]}
    # one survivor holds the whole image: it provides every stripe the others lack
    got = plan_providers([f("a", hbm=(5, 0, total), shm=5), f("a")], 5, total, chunk)
    assert got == {"a": [(0, 6 * MB, 0, False), (6 * MB, total, 0, False)]}
    # striped save, then rank 1 was replaced: it re-stages its own stripe from shm
    got = plan_providers([f("a", hbm=(5, 0, 6 * MB), shm=5), f("a", shm=5)], 5, total, chunk)
    assert got == {"a": [(0, 6 * MB, 0, False), (6 * MB, total, 1, True)]}
    # a stale HBM image (older step) does not count
    got = plan_providers([f("a", hbm=(4, 0, total), shm=5), f("a", shm=5)], 5, total, chunk)
    assert got["a"][0][3] is True and got["a"][1][3] is True
    # a host with neither HBM nor shm for a stripe: not handled
    assert plan_providers([f("a", hbm=(5, 0, total)), f("b")], 5, total, chunk) is None
    # two hosts, each stripes its own shm; more ranks than chunks leaves empty stripes
    got = plan_providers([f("a", shm=5), f("b", shm=5)], 5, 100, chunk)
    assert got == {"a": [(0, 100, 0, True)], "b": [(0, 100, 1, True)]}
    got = plan_providers([f("a", shm=5)] * 4, 5, 3 * MB, chunk)
    assert [s[:2] for s in got["a"]] == [(0, 2 * MB), (2 * MB, 3 * MB), (3 * MB, 3 * MB),
                                         (3 * MB, 3 * MB)]


# ------------------------------------------------------- DMA turn-taking
def test_dma_gate_slots(monkeypatch):
    import threading
    import time
    from dlrover_wuqiong_b200.transfer import DmaGate, default_bounce_workers
    monkeypatch.setenv("FCP_DMA_WRITERS", "2")
    assert not DmaGate(0, 2).enabled and not DmaGate(1, 1).enabled  # few ranks: no pacing
    a, b, c = DmaGate(0, 4), DmaGate(2, 4), DmaGate(1, 4)  # ranks 0 and 2 share slot 0
    assert a.enabled and b.enabled and c.enabled
    a.acquire()
    c.acquire()  # another slot: does not wait
    got = []
    t = threading.Thread(target=lambda: (b.acquire(), got.append(time.perf_counter()), b.release()))
    t0 = time.perf_counter()
    t.start()
    time.sleep(0.3)
    assert not got  # rank 2 waits for rank 0's turn to end
    a.release()
    t.join(5)
    assert got and got[0] - t0 >= 0.3 and b.wait_s >= 0.25
    c.release()
    for g in (a, b, c):
        g.close()
    monkeypatch.setenv("FCP_DMA_WRITERS", "0")
    assert not DmaGate(0, 8).enabled
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "8")
    monkeypatch.delenv("FCP_BOUNCE_WORKERS", raising=False)
    assert 2 <= default_bounce_workers() <= 8
    monkeypatch.setenv("FCP_BOUNCE_WORKERS", "5")
    assert default_bounce_workers() == 5
