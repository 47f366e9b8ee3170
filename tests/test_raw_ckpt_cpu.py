"""CPU: raw-image persistence.  The agent writes the shm bytes
and the meta subtree instead of torch.save-ing views; the files must hold the
same state the reference's saver would have persisted
(dlrover/python/elastic_agent/torch/ckpt_saver.py:914-921) and every loader
must accept either format."""
import argparse
import os
import subprocess
import sys
import tempfile
import time

import pytest
import torch
import torch.nn as nn

os.environ.setdefault("ROLE_NAME", "dlrover-trainer")
os.environ.setdefault("TORCHELASTIC_RUN_ID", "fcp_test_%d" % os.getpid())

from dlrover_wuqiong_b200 import plan, raw_ckpt  # noqa: E402
from dlrover_wuqiong_b200.ckpt_saver import AsyncCheckpointSaver, _host_copy_leaves  # noqa: E402
from dlrover_wuqiong_b200.common.constants import CheckpointConstant, NodeEnv  # noqa: E402
from dlrover_wuqiong_b200.common.multi_process import clear_sock_dir  # noqa: E402
from dlrover_wuqiong_b200.flash_checkpoint.ddp import DdpCheckpointer, StorageType  # noqa: E402


def _same(a, b):
    if isinstance(a, dict):
        assert isinstance(b, dict) and list(a.keys()) == list(b.keys())
        for k in a:
            _same(a[k], b[k])
    elif isinstance(a, list):
        assert isinstance(b, list) and len(a) == len(b)
        for x, y in zip(a, b):
            _same(x, y)
    elif torch.is_tensor(a):
        assert torch.is_tensor(b) and a.dtype == b.dtype
        if a.numel() == 0:
            assert b.numel() == 0  # empty tensors come back as shape (0,), as in the reference
        else:
            assert a.shape == b.shape
            assert a.contiguous().reshape(-1).view(torch.uint8).equal(
                b.contiguous().reshape(-1).view(torch.uint8))
    else:
        assert a == b


def _state():
    g = torch.Generator().manual_seed(3)
    return {
        "model_states": {
            "model": {"w": torch.randn(33, 7, generator=g), "b": torch.randn(5, generator=g).bfloat16(),
                      "empty": torch.empty(0, 4), "scalar": torch.tensor(2.5, dtype=torch.float64)},
            "args": argparse.Namespace(lr=0.1, name="x"),
            "rng": [torch.arange(17, dtype=torch.uint8), (1, 2, "tuple-leaf")],
            "step": 12,
        },
        "optim_states": {
            "state": {0: {"exp_avg": torch.randn(33, 7, generator=g).t(),  # non-contiguous
                          "step": torch.tensor(3.0)}},
            "param_groups": [{"lr": 0.1, "params": [0]}],
        },
    }


def _image_of(state):
    meta, leaves, total = plan.build_layout(state)
    buf = bytearray(total)
    _host_copy_leaves(leaves, buf)
    return meta, buf, total


def test_each_state_name_round_trips_and_converts(tmp_path):
    state = _state()
    meta, buf, total = _image_of(state)
    lo_m, hi_m = raw_ckpt.subtree_range(meta["model_states"])
    lo_o, hi_o = raw_ckpt.subtree_range(meta["optim_states"])
    assert lo_m == 0 and hi_m == lo_o and hi_o == total  # contiguous, in traversal order
    for name in state:
        path = str(tmp_path / (name + ".raw"))
        raw_ckpt.write_raw(raw_ckpt.RawImage(meta[name], buf, step=12), path)
        assert raw_ckpt.is_raw(path)
        header, data_off = raw_ckpt.read_header(path)
        assert data_off % raw_ckpt.DATA_ALIGN == 0 and header["step"] == 12
        assert os.path.getsize(path) == data_off + header["nbytes"]
        back = raw_ckpt.read_raw(path)
        _same(state[name], back)
        _same(state[name], raw_ckpt.load_any(path))
        # tensors are private and writable
        next(v for v in back.values() if isinstance(v, dict)).popitem()
        pt = str(tmp_path / (name + ".pt"))
        raw_ckpt.convert_to_pt(path, pt)
        assert not raw_ckpt.is_raw(pt)
        _same(state[name], torch.load(pt, weights_only=False))
        _same(state[name], raw_ckpt.load_any(pt))
    # the command-line converter
    out = str(tmp_path / "cli.pt")
    subprocess.check_call([sys.executable, "-m", "dlrover_wuqiong_b200.raw_ckpt",
                           str(tmp_path / "optim_states.raw"), out],
                          cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    _same(state["optim_states"], torch.load(out, weights_only=False))


def test_no_tensor_state_and_truncated_file(tmp_path):
    state = {"model_states": {"step": 3, "empty": torch.empty(0)}}
    meta, buf, total = _image_of(state)
    assert total == 0
    path = str(tmp_path / "s.raw")
    raw_ckpt.write_raw(raw_ckpt.RawImage(meta["model_states"], buf), path)
    _same(state["model_states"], raw_ckpt.read_raw(path))
    big = {"model_states": {"w": torch.arange(5000, dtype=torch.int32)}}
    meta, buf, total = _image_of(big)
    raw_ckpt.write_raw(raw_ckpt.RawImage(meta["model_states"], buf), path)
    with open(path, "r+b") as f:
        f.truncate(os.path.getsize(path) - 100)
    with pytest.raises(ValueError):
        raw_ckpt.read_raw(path)
    with open(path, "wb") as f:
        f.write(b"not an image")
    assert not raw_ckpt.is_raw(path)
    with pytest.raises(ValueError):
        raw_ckpt.read_header(path)


def test_payload_larger_than_one_write_block(tmp_path, monkeypatch):
    monkeypatch.setattr(raw_ckpt, "_WRITE_BLOCK", 1 << 16)  # 5 blocks + a ragged one
    g = torch.Generator().manual_seed(9)
    state = {"model_states": {"a": torch.randint(0, 255, (5 * 65536 + 4321,), dtype=torch.uint8,
                                                 generator=g), "b": torch.randn(11, generator=g)}}
    meta, buf, total = _image_of(state)
    path = str(tmp_path / "big.raw")
    raw_ckpt.write_raw(raw_ckpt.RawImage(meta["model_states"], buf), path)
    _same(state["model_states"], raw_ckpt.read_raw(path))


def test_bad_format_name(monkeypatch):
    monkeypatch.setenv(raw_ckpt.PERSIST_FORMAT_ENV, "zip")
    with pytest.raises(ValueError):
        raw_ckpt.persist_format()


class SimpleNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.fc1 = nn.Linear(64, 32)
        self.fc2 = nn.Linear(32, 10)


@pytest.fixture
def saver_env(monkeypatch):
    monkeypatch.setenv(NodeEnv.NODE_NUM, "1")
    monkeypatch.setenv(NodeEnv.NODE_RANK, "0")
    monkeypatch.setenv(raw_ckpt.PERSIST_FORMAT_ENV, "raw")
    AsyncCheckpointSaver._saver_instance = None
    AsyncCheckpointSaver.start_async_saving_ckpt()
    yield
    if AsyncCheckpointSaver._saver_instance:
        AsyncCheckpointSaver._saver_instance.close()
        AsyncCheckpointSaver._saver_instance = None
    clear_sock_dir()


def _wait_for(pred, timeout=20):
    t0 = time.time()
    while time.time() - t0 < timeout:
        if pred():
            return True
        time.sleep(0.1)
    return False


def test_agent_persists_raw_and_checkpointer_loads_it(saver_env):
    model = SimpleNet()
    opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)
    model.fc2(model.fc1(torch.randn(4, 64))).sum().backward()
    opt.step()
    with tempfile.TemporaryDirectory() as tmpdir:
        ckpt = DdpCheckpointer(tmpdir)
        sd = {"model": model.state_dict(), "optimizer": opt.state_dict(), "step": 200}
        assert ckpt.save_checkpoint(200, sd, storage_type=StorageType.DISK)
        tracker = os.path.join(tmpdir, CheckpointConstant.TRACER_FILE_NAME)
        assert _wait_for(lambda: os.path.exists(tracker) and open(tracker).read() == "200")
        path = os.path.join(tmpdir, "200", "rank_0.pt")
        assert raw_ckpt.is_raw(path)
        header, data_off = raw_ckpt.read_header(path)
        assert header["step"] == 200 and header["image"]["offset"] == 0
        assert header["nbytes"] == ckpt._engine._shm_handler._buffer_size
        # payload == the shm image, byte for byte
        with open(path, "rb") as f:
            f.seek(data_off)
            assert f.read() == bytes(ckpt._engine._shm_handler.shared_memory.buf)
        _same(sd, raw_ckpt.read_raw(path))
        # memory gone -> the checkpointer falls back to storage and reads the raw file
        ckpt._engine._shm_handler.metadata.set({})
        from_disk = ckpt.load_checkpoint()
        _same(sd, from_disk)
        _same(sd, ckpt.load_checkpoint(path))
        # ... and in place, from the raw file and from a torch.save file
        pt = os.path.join(tmpdir, "as_pt.pt")
        raw_ckpt.convert_to_pt(path, pt)
        for source in (path, pt):
            into = {"model": {k: torch.zeros_like(v) for k, v in model.state_dict().items()},
                    "optimizer": {"state": {i: {"momentum_buffer": torch.zeros_like(p)}
                                            for i, p in enumerate(model.parameters())},
                                  "param_groups": opt.state_dict()["param_groups"]},
                    "step": 0}
            live = into["model"]["fc1.weight"]
            back = ckpt.load_checkpoint(source, into=into)
            _same(sd, back)
            assert back["model"]["fc1.weight"] is live and torch.equal(live, model.fc1.weight)
        with pytest.raises(ValueError):
            ckpt.load_checkpoint(path, into={"model": {}, "step": 0})
        ckpt.close()
