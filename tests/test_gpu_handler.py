"""GPU parity of the reference-facing API: SharedMemoryHandler / DdpCheckpointer
with CUDA state dicts must leave exactly the reference's bytes in shm (golden
vectors + oracle) and restore bit-exact tensors."""
import os
import tempfile

import numpy as np
import pytest
import torch
import torch.nn as nn

os.environ.setdefault("ROLE_NAME", "dlrover-trainer")
os.environ.setdefault("TORCHELASTIC_RUN_ID", "fcp_test_%d" % os.getpid())

from dlrover_wuqiong_b200 import _native  # noqa: E402
from dlrover_wuqiong_b200.ckpt_saver import (  # noqa: E402
    DLROVER_CKPT_CONFIG_KEY,
    FCP_CHECKSUM_KEY,
    AsyncCheckpointSaver,
    CheckpointConfig,
    SharedMemoryHandler,
)
from dlrover_wuqiong_b200.common.constants import NodeEnv  # noqa: E402
from dlrover_wuqiong_b200.common.multi_process import clear_sock_dir  # noqa: E402
from dlrover_wuqiong_b200.flash_checkpoint.ddp import DdpCheckpointer, StorageType  # noqa: E402
from oracle import oracle  # noqa: E402

pytestmark = pytest.mark.gpu


def _to_dev(sd, keep_cpu=()):
    def mv(v):
        return v.cuda() if torch.is_tensor(v) else v
    out = oracle.traverse(sd, mv)
    for k in keep_cpu:
        out[k] = sd[k]
    return out


@pytest.mark.parametrize("name", ["simplenet_sgd_9640", "toymodel_adam_1632", "quirks",
                                  "mlp_adamw_stepped", "noncontig"])
def test_handler_cuda_state_dict_matches_reference_bytes(golden, name):
    case = golden["cases"][name]
    h = SharedMemoryHandler(41, host=True)
    try:
        sd = _to_dev(case["state_dict"])
        if name == "noncontig":  # keep real strided device views
            base = case["state_dict"]["t"].t().cuda()
            sd["col_slice"], sd["row_step"], sd["t"] = base[:, 3:11], base[::3], base.t()
        sd[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=5, paths={})
        h.save_state_dict(sd)
        assert h.shared_memory.size == case["size"]
        assert bytes(h.shared_memory.buf[:case["size"]]) == case["shm"]
        meta = h.metadata.get()
        sums = meta[FCP_CHECKSUM_KEY]["sums"]
        want = oracle.chunk_crc32c(np.frombuffer(case["shm"], dtype=np.uint8),
                                   _native.DEFAULT_CHUNK_BYTES)
        assert sums == want.tolist()
        # reference-style load: CPU views with the reference's values
        out = h.load_state_dict()
        flat_in, _, _ = oracle.flatten(case["state_dict"])
        flat_out, _, _ = oracle.flatten({k: v for k, v in out.items()
                                         if k != DLROVER_CKPT_CONFIG_KEY})
        for (a, _), (b, _) in zip(flat_in, flat_out):
            assert oracle.tensor_bytes(a).tobytes() == oracle.tensor_bytes(b).tobytes()
        # device restore into fresh tensors
        fresh = h.load_state_dict(device="cuda")
        flat_f, _, _ = oracle.flatten({k: v for k, v in fresh.items()
                                       if k != DLROVER_CKPT_CONFIG_KEY})
        for (a, _), (b, _) in zip(flat_in, flat_f):
            if a.numel():
                assert b.is_cuda
                assert oracle.tensor_bytes(a).tobytes() == oracle.tensor_bytes(b).tobytes()
        if "second" in case:
            sd2 = _to_dev(case["second"]["state_dict"])
            sd2[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=6, paths={})
            h.save_state_dict(sd2)
            assert bytes(h.shared_memory.buf[:case["size"]]) == case["second"]["shm"]
    finally:
        h.unlink()
        h.close()


def test_mixed_host_and_device_leaves_and_corruption():
    """CPU-resident leaves (AdamW's default `step`, rng states) ride through
    the arena; a flipped bit in shm is caught by the CRC on restore."""
    torch.manual_seed(3)
    model = nn.Sequential(nn.Linear(301, 257), nn.Linear(257, 33)).cuda()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
    model(torch.randn(8, 301, device="cuda")).sum().backward()
    opt.step()
    osd = opt.state_dict()
    assert not osd["state"][0]["step"].is_cuda  # torch keeps it on the host
    sd = {"model": model.state_dict(), "optimizer": osd,
          "rng": [torch.get_rng_state(), torch.cuda.get_rng_state()], "step": 1,
          DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=1, paths={})}
    _, total, expect = oracle.pack({k: v for k, v in sd.items() if k != DLROVER_CKPT_CONFIG_KEY})
    h = SharedMemoryHandler(42, host=True)
    try:
        h.save_state_dict(sd)
        assert bytes(h.shared_memory.buf[:total]) == expect.tobytes()
        into = {"model": {k: torch.zeros_like(v) for k, v in model.state_dict().items()},
                "optimizer": {"state": {i: {k: torch.zeros_like(v) for k, v in s.items()}
                                        for i, s in osd["state"].items()},
                              "param_groups": osd["param_groups"]},
                "rng": [torch.zeros_like(sd["rng"][0]), torch.zeros_like(sd["rng"][1])],
                "step": 0}
        out = h.load_state_dict(into=into)
        assert out["step"] == 1
        for k, v in model.state_dict().items():
            assert torch.equal(into["model"][k], v)
        for i, s in osd["state"].items():
            for k, v in s.items():
                assert torch.equal(into["optimizer"]["state"][i][k], v)
        assert torch.equal(into["rng"][0], sd["rng"][0])
        h.shared_memory.buf[total // 3] ^= 0x10
        with pytest.raises(_native.ChecksumMismatch):
            h.load_state_dict(into=into)
    finally:
        h.unlink()
        h.close()


def test_ddp_checkpointer_gpu_roundtrip_and_async():
    os.environ[NodeEnv.NODE_NUM] = "1"
    os.environ[NodeEnv.NODE_RANK] = "0"
    AsyncCheckpointSaver._saver_instance = None
    AsyncCheckpointSaver.start_async_saving_ckpt()
    try:
        torch.manual_seed(9)
        model = nn.Sequential(nn.Linear(512, 1024), nn.Linear(1024, 77)).cuda()
        opt = torch.optim.AdamW(model.parameters(), lr=1e-3, fused=True)
        model(torch.randn(4, 512, device="cuda")).sum().backward()
        opt.step()
        with tempfile.TemporaryDirectory() as tmpdir:
            for async_save in (False, True):
                ckpt = DdpCheckpointer(tmpdir, async_save=async_save)
                sd = {"model": model.state_dict(), "optimizer": opt.state_dict(), "step": 10}
                meta, total, expect = oracle.pack(sd)
                ref = oracle.unpack(meta, expect)  # snapshot before any mutation
                assert ckpt.save_checkpoint(10, sd, storage_type=StorageType.MEMORY)
                if async_save:
                    # the tensors may change as soon as save returns: the image
                    # in shm must still be the snapshot
                    with torch.no_grad():
                        for p in model.parameters():
                            p.add_(1.0)
                    ckpt.wait_latest_checkpoint()
                shm = ckpt._engine._shm_handler.shared_memory
                assert bytes(shm.buf[:total]) == expect.tobytes()
                back = ckpt.load_checkpoint(device="cuda")
                assert back["step"] == 10
                for k in sd["model"]:
                    assert back["model"][k].is_cuda
                    assert torch.equal(back["model"][k].cpu(), ref["model"][k])
                ckpt.close()
    finally:
        if AsyncCheckpointSaver._saver_instance:
            AsyncCheckpointSaver._saver_instance.close()
            AsyncCheckpointSaver._saver_instance = None
        clear_sock_dir()


def test_cold_restore_of_unpinned_segment_uses_bounce_ring():
    """A restarted trainer attaches to a segment it has not pinned: the image
    is streamed through the pinned ring (no cudaHostRegister on the critical
    path), verified and unpacked; the segment is pinned in the background."""
    torch.manual_seed(21)
    tensors = {"w%d" % i: torch.randn(1 << 20, device="cuda") for i in range(40)}  # 160 MiB
    tensors["odd"] = torch.randn(12345, device="cuda").to(torch.bfloat16)
    sd = dict(tensors)
    sd["step"] = 4
    sd[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=4, paths={})
    writer = SharedMemoryHandler(43, host=True)
    try:
        writer.save_state_dict(sd)
        writer._drop_engine()  # as if the writing process were gone
        reader = SharedMemoryHandler(43, host=False)
        into = {k: torch.zeros_like(v) for k, v in tensors.items()}
        into["step"] = 0
        out = reader.load_state_dict(into=into)
        assert reader.last_stats["restore_path"] == "bounced"
        assert out["step"] == 4
        for k, v in tensors.items():
            assert torch.equal(into[k], v), k
        # second restore in the same process: pinned by now
        reader.load_state_dict(into=into)
        assert reader.last_stats["restore_path"] == "registered"
        reader._drop_engine()
        writer.shared_memory.buf[5 << 20] ^= 0x01
        with pytest.raises(_native.ChecksumMismatch):
            reader.load_state_dict(into=into)
        reader.close()
    finally:
        writer.unlink()
        writer.close()


def test_optional_xxh64_is_recorded_and_verified(monkeypatch):
    monkeypatch.setenv("FCP_XXH64", "1")
    torch.manual_seed(31)
    tensors = {"w%d" % i: torch.randn(700001 + i, device="cuda") for i in range(6)}
    sd = dict(tensors)
    sd[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=2, paths={})
    h = SharedMemoryHandler(44, host=True)
    try:
        h.save_state_dict(sd)
        entry = h.metadata.get()[FCP_CHECKSUM_KEY]
        total = h.shared_memory.size
        image = np.frombuffer(h.shared_memory.buf, dtype=np.uint8)[:total]
        chunk = entry["chunk_bytes"]
        want = [oracle.xxh64(image[i * chunk:(i + 1) * chunk])
                for i in range((total + chunk - 1) // chunk)]
        assert entry["xxh64"] == want
        del image
        into = {k: torch.zeros_like(v) for k, v in tensors.items()}
        h.load_state_dict(into=into)
        for k, v in tensors.items():
            assert torch.equal(into[k], v)
        # a wrong recorded hash is noticed even though the CRCs agree
        meta = h.metadata.get()
        meta[FCP_CHECKSUM_KEY]["xxh64"][0] ^= 1
        h.metadata.set(meta)
        with pytest.raises(_native.ChecksumMismatch):
            h.load_state_dict(into=into)
    finally:
        h.unlink()
        h.close()


def test_raw_persisted_file_restores_in_place_through_the_scatter_kernel(monkeypatch):
    """DLROVER_FCP_PERSIST_FORMAT=raw: the agent writes the shm image itself;
    a trainer whose memory checkpoint is gone restores that file straight into
    its CUDA tensors (file mapping -> pinned ring -> scatter kernel, CRC
    verified), and the same file still reads as the reference's state dict."""
    import time

    from dlrover_wuqiong_b200 import raw_ckpt
    from dlrover_wuqiong_b200.common.constants import CheckpointConstant

    monkeypatch.setenv(NodeEnv.NODE_NUM, "1")
    monkeypatch.setenv(NodeEnv.NODE_RANK, "0")
    monkeypatch.setenv(raw_ckpt.PERSIST_FORMAT_ENV, "raw")
    AsyncCheckpointSaver._saver_instance = None
    AsyncCheckpointSaver.start_async_saving_ckpt()
    try:
        torch.manual_seed(17)
        model = nn.Sequential(nn.Linear(1024, 2048), nn.Linear(2048, 1031)).cuda()
        opt = torch.optim.AdamW(model.parameters(), lr=1e-3, fused=True)
        model(torch.randn(4, 1024, device="cuda")).sum().backward()
        opt.step()
        with tempfile.TemporaryDirectory() as tmpdir:
            ckpt = DdpCheckpointer(tmpdir)
            sd = {"model": model.state_dict(), "optimizer": opt.state_dict(), "step": 30}
            meta, total, expect = oracle.pack(sd)
            ref = oracle.unpack(meta, expect)
            assert ckpt.save_checkpoint(30, sd, storage_type=StorageType.DISK)
            tracker = os.path.join(tmpdir, CheckpointConstant.TRACER_FILE_NAME)
            t0 = time.time()
            while not (os.path.exists(tracker) and open(tracker).read() == "30"):
                assert time.time() - t0 < 30, "the agent did not persist step 30"
                time.sleep(0.1)
            path = os.path.join(tmpdir, "30", "rank_0.pt")
            header, data_off = raw_ckpt.read_header(path)
            with open(path, "rb") as f:
                f.seek(data_off)
                assert f.read() == expect.tobytes()  # the reference's bytes
            # memory checkpoint gone -> storage, in place
            ckpt._engine._shm_handler.metadata.set({})
            into = {"model": {k: torch.zeros_like(v) for k, v in sd["model"].items()},
                    "optimizer": {"state": {i: {k: torch.zeros_like(v) for k, v in st.items()}
                                            for i, st in sd["optimizer"]["state"].items()},
                                  "param_groups": sd["optimizer"]["param_groups"]},
                    "step": 0}
            back = ckpt.load_checkpoint(into=into)
            assert ckpt._engine._shm_handler.last_stats["restore_path"] == "bounced_file"
            assert back["step"] == 30
            for k in sd["model"]:
                assert into["model"][k].is_cuda
                assert torch.equal(into["model"][k].cpu(), ref["model"][k]), k
            for i, st in ref["optimizer"]["state"].items():
                for k, v in st.items():
                    assert torch.equal(into["optimizer"]["state"][i][k].cpu(), v), (i, k)
            # one flipped payload bit on disk is caught by the kernel's CRC
            with open(path, "r+b") as f:
                f.seek(data_off + total // 2)
                b = f.read(1)
                f.seek(data_off + total // 2)
                f.write(bytes([b[0] ^ 0x10]))
            with pytest.raises(_native.ChecksumMismatch):
                ckpt.load_checkpoint(path, into=into)
            ckpt.close()
    finally:
        if AsyncCheckpointSaver._saver_instance:
            AsyncCheckpointSaver._saver_instance.close()
            AsyncCheckpointSaver._saver_instance = None
        clear_sock_dir()


def test_save_leg_is_probed_and_chosen_at_run_time(monkeypatch):
    """Force the tuner to distrust the copy engine: the third save must go
    through the copy kernel, every image must still be the reference's bytes,
    and the process must settle on one of the two legs."""
    from dlrover_wuqiong_b200 import transfer

    monkeypatch.delenv("FCP_TRANSFER_MODE", raising=False)
    monkeypatch.setattr(transfer.SaveLegTuner, "GOOD_GBPS", 1e9)
    monkeypatch.setattr(transfer.SaveLegTuner, "MIN_BYTES", 1 << 20)
    torch.manual_seed(5)
    tensors = {"w%d" % i: torch.randn((1 << 20) + 3 * i, device="cuda") for i in range(24)}
    sd = dict(tensors)
    sd[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=1, paths={})
    h = SharedMemoryHandler(45, host=True)
    try:
        modes = []
        for step in range(5):
            with torch.no_grad():
                tensors["w3"].add_(1.0)
            meta, total, expect = oracle.pack({k: v for k, v in sd.items()
                                               if k != DLROVER_CKPT_CONFIG_KEY})
            h.save_state_dict(sd)
            modes.append(h._engine.stats["save_mode"])
            assert bytes(h.shared_memory.buf[:total]) == expect.tobytes(), step
            sums = h.metadata.get()[FCP_CHECKSUM_KEY]["sums"]
            assert sums == oracle.chunk_crc32c(expect, _native.DEFAULT_CHUNK_BYTES).tolist()
        assert modes[:3] == [_native.MODE_STAGED, _native.MODE_STAGED, _native.MODE_STAGED_KERNEL]
        tuner = h._engine.tuner
        assert tuner.settled and set(tuner.rates) == {"engine", "kernel"}
        assert modes[3] == modes[4] == tuner.choice
        into = {k: torch.zeros_like(v) for k, v in tensors.items()}
        h.load_state_dict(into=into)
        for k, v in tensors.items():
            assert torch.equal(into[k], v)
    finally:
        h.unlink()
        h.close()


def _random_nested(rng, depth=0):
    """A random nested state dict: dicts / lists as containers, CUDA and host
    tensors of mixed dtypes (0-dim, empty, strided views, odd sizes), tuples
    and scalars as opaque leaves — the shapes the reference's traversal meets
    (SAVER:94-109)."""
    dtypes = [torch.float32, torch.bfloat16, torch.float16, torch.float64, torch.int64,
              torch.int32, torch.int16, torch.uint8, torch.bool]

    def tensor():
        dt = dtypes[int(rng.integers(len(dtypes)))]
        kind = int(rng.integers(8))
        if kind == 0:
            shape = ()
        elif kind == 1:
            shape = (0, int(rng.integers(1, 5)))
        elif kind == 2:
            shape = (int(rng.integers(1, 40)), int(rng.integers(1, 40)), int(rng.integers(1, 7)))
        else:
            shape = (int(rng.integers(1, 30000)),)
        n = 1
        for s in shape:
            n *= s
        raw = torch.from_numpy(rng.integers(0, 256, max(n, 1) * 8, dtype=np.uint8))
        if dt == torch.bool:
            t = (raw[:n] & 1).bool().reshape(shape)
        else:
            t = raw[:n * torch.empty((), dtype=dt).element_size()].view(dt).reshape(shape)
        if dt.is_floating_point:
            t = torch.nan_to_num(t.float(), nan=1.0, posinf=2.0, neginf=-2.0).to(dt)
        if int(rng.integers(5)) != 0:
            t = t.cuda()
        if len(shape) == 3 and int(rng.integers(2)):
            t = t.transpose(0, 1)  # strided view: must be linearised row-major
        return t

    def leaf():
        k = int(rng.integers(10))
        if k < 6:
            return tensor()
        return [3, 2.5, "s", None, (1, torch.ones(2)), True][int(rng.integers(6))]

    if depth >= 3:
        return leaf()
    out = {}
    for i in range(int(rng.integers(1, 6))):
        k = int(rng.integers(4))
        key = ["k%d" % i, i, "layer.%d.weight" % i][int(rng.integers(3))]
        if k == 0:
            out[key] = _random_nested(rng, depth + 1)
        elif k == 1:
            out[key] = [leaf() if int(rng.integers(3)) else _random_nested(rng, depth + 1)
                        for _ in range(int(rng.integers(0, 4)))]
        else:
            out[key] = leaf()
    return out


@pytest.mark.parametrize("seed", range(12))
def test_random_nested_state_dicts_match_the_oracle(seed):
    rng = np.random.default_rng(1000 + seed)
    sd = {"model_states": _random_nested(rng), "optim_states": _random_nested(rng)}
    flat, _, total = oracle.flatten(sd)
    if not any(t.is_cuda and t.numel() for t, _ in flat):
        sd["model_states"]["force_gpu"] = torch.arange(7, device="cuda", dtype=torch.int16)
    meta, total, expect = oracle.pack(sd)
    sd[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=seed + 1, paths={})
    h = SharedMemoryHandler(50 + seed, host=True)
    try:
        h.save_state_dict(sd)
        assert h.shared_memory.size == total
        assert bytes(h.shared_memory.buf[:total]) == expect.tobytes()
        sums = h.metadata.get()[FCP_CHECKSUM_KEY]["sums"]
        assert sums == oracle.chunk_crc32c(expect, _native.DEFAULT_CHUNK_BYTES).tolist()
        # reference-style views and a device restore both reproduce the oracle's unpack
        want = oracle.unpack(meta, expect)
        got = h.load_state_dict()
        fresh = h.load_state_dict(device="cuda")
        fw, _, _ = oracle.flatten(want)
        for other in (got, fresh):
            fo, _, _ = oracle.flatten({k: v for k, v in other.items()
                                       if k != DLROVER_CKPT_CONFIG_KEY})
            assert len(fo) == len(fw)
            for (a, _), (b, _) in zip(fw, fo):
                assert a.dtype == b.dtype
                assert oracle.tensor_bytes(a).tobytes() == oracle.tensor_bytes(b).tobytes()
    finally:
        h.unlink()
        h.close()
