"""CPU: host-side mirror of the reference interface — IPC primitives, the shm
handler against the reference's golden bytes, the agent saver and the DDP
checkpointer.  Shaped after the reference's own tests
(dlrover/python/tests/test_multi_process.py, test_ckpt_saver.py,
dlrover/trainer/tests/torch/checkpoint_egine_test.py, ddp_checkpointer_test.py)."""
import os
import tempfile
import time
from pathlib import Path

import pytest
import torch
import torch.nn as nn

os.environ.setdefault("ROLE_NAME", "dlrover-trainer")
os.environ.setdefault("TORCHELASTIC_RUN_ID", "fcp_test_%d" % os.getpid())

from dlrover_wuqiong_b200.ckpt_saver import (  # noqa: E402
    DLROVER_CKPT_CONFIG_KEY,
    AsyncCheckpointSaver,
    CheckpointConfig,
    CheckpointEvent,
    CheckpointEventType,
    DdpCheckpointSaver,
    SharedMemoryHandler,
    TempDirCheckpointSaver,
    TensorMeta,
)
from dlrover_wuqiong_b200.common.constants import CheckpointConstant, NodeEnv  # noqa: E402
from dlrover_wuqiong_b200.common.multi_process import (  # noqa: E402
    SharedDict,
    SharedLock,
    SharedMemory,
    SharedQueue,
    clear_sock_dir,
)
from dlrover_wuqiong_b200.common.storage import (  # noqa: E402
    KeepLatestStepStrategy,
    KeepStepIntervalStrategy,
    PosixDiskStorage,
    PosixStorageWithDeletion,
)
from dlrover_wuqiong_b200.flash_checkpoint.ddp import DdpCheckpointer, StorageType  # noqa: E402
from dlrover_wuqiong_b200.flash_checkpoint.full_ckpt_engine import FullCheckpointEngine  # noqa: E402


class SimpleNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.fc1 = nn.Linear(64, 32)
        self.fc2 = nn.Linear(32, 10)


# ------------------------------------------------------------------ IPC
def test_shared_lock():
    server = SharedLock("t_lock", create=True)
    client = SharedLock("t_lock", create=False)
    assert client.acquire()
    assert server.locked()
    assert client.locked() is True
    assert not client.acquire(blocking=False)
    client.release()
    assert not server.locked()
    assert client.locked() is False
    client.release()  # releasing an unlocked lock is harmless
    server.unlink()


def test_shared_queue():
    server = SharedQueue("t_queue", create=True)
    client = SharedQueue("t_queue", create=False)
    assert client.empty()
    client.put({"a": [1, 2, 3]})
    assert server.qsize() == 1 and client.qsize() == 1
    assert not client.empty()
    assert server.get() == {"a": [1, 2, 3]}
    server.unlink()


def test_shared_dict_large_message():
    server = SharedDict("t_dict", create=True)
    client = SharedDict("t_dict", create=False)
    big = {"k%d" % i: TensorMeta((i, 3), torch.float32, 4, 3 * i, 12 * i) for i in range(20000)}
    client.set(big)  # > 1 MiB pickled: exercises the sized receive
    assert server.get() == big
    assert client.get() == big
    assert client.get(local=True) == big
    server.unlink()


def test_shared_memory_survives_close():
    name = "fcp_t_shm_%d" % os.getpid()
    a = SharedMemory(name=name, create=True, size=1024)
    a.buf[0:4] = b"abcd"
    a.close()
    b = SharedMemory(name=name)  # still there: close() does not unlink
    assert bytes(b.buf[0:4]) == b"abcd" and b.size == 1024
    b.close()
    b.unlink()
    with pytest.raises(FileNotFoundError):
        SharedMemory(name=name)


# ------------------------------------------------------------ shm handler
@pytest.mark.parametrize("name", ["simplenet_sgd_9640", "toymodel_adam_1632", "quirks",
                                  "mlp_adamw_stepped", "noncontig"])
def test_handler_writes_reference_bytes(golden, name):
    case = golden["cases"][name]
    h = SharedMemoryHandler(31, host=True)
    try:
        sd = dict(case["state_dict"])
        sd[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=5, paths={})
        h.save_state_dict(sd)
        assert h.shared_memory.size == case["size"]
        assert bytes(h.shared_memory.buf[:case["size"]]) == case["shm"]
        meta = h.metadata.get()
        assert meta[DLROVER_CKPT_CONFIG_KEY].writing_shm is False
        assert meta[DLROVER_CKPT_CONFIG_KEY].step == 5

        def check(ref, mine):
            if isinstance(ref, dict) and ref.get("__tensor_meta__"):
                assert isinstance(mine, TensorMeta)
                assert (tuple(ref["shape"]), ref["dtype"], ref["element_size"], ref["numel"],
                        ref["offset"]) == (tuple(mine.shape), mine.dtype, mine.element_size,
                                           mine.numel, mine.offset)
            elif isinstance(ref, dict) and ref.get("__ckpt_config__"):
                assert isinstance(mine, CheckpointConfig)
            elif isinstance(ref, dict):
                assert list(ref.keys()) == list(mine.keys())
                for k in ref:
                    check(ref[k], mine[k])
            elif isinstance(ref, list):
                assert len(ref) == len(mine)
                for a, b in zip(ref, mine):
                    check(a, b)
            elif not isinstance(ref, tuple):
                assert ref == mine

        check(case["meta"], meta)
        out = h.load_state_dict()
        if name == "simplenet_sgd_9640":
            for k, v in case["state_dict"]["model_states"].items():
                assert torch.equal(out["model_states"][k], v)
        if "second" in case:
            sd2 = dict(case["second"]["state_dict"])
            sd2[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=6, paths={})
            h.save_state_dict(sd2)
            assert bytes(h.shared_memory.buf[:case["size"]]) == case["second"]["shm"]
            assert h.metadata.get()["step"] == 2
    finally:
        h.unlink()
        h.close()


def test_handler_tensor_meta_and_resize():
    # dlrover/python/tests/test_ckpt_saver.py:85-92, 165-174
    h = SharedMemoryHandler(32, host=True)
    try:
        m = h._create_tensor_meta(torch.rand(10, 10))
        assert (m.shape, m.dtype, m.element_size, m.numel, m.offset) == (
            (10, 10), torch.float32, 4, 100, 0)
        h._buffer_size = 0
        sd = {"w": torch.rand(4, 4), DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=1)}
        h.save_state_dict(sd)
        assert h.shared_memory.size == 64
        sd = {"w": torch.rand(8, 4), DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=2)}
        h.save_state_dict(sd)  # structure changed: detected, segment re-created
        assert h.shared_memory.size == 128
        assert torch.equal(h.load_state_dict()["w"], sd["w"])
    finally:
        h.unlink()
        h.close()


def test_handler_load_into_and_empty_states():
    h = SharedMemoryHandler(33, host=True)
    try:
        assert h.load_state_dict() == {}
        assert h.no_checkpint_state()
        model = SimpleNet()
        sd = {"model": model.state_dict(), "step": 7, "lst": [torch.arange(5), "s"],
              DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=7)}
        h.save_state_dict(sd)
        assert not h.no_checkpint_state()
        into = {"model": {k: torch.zeros_like(v) for k, v in model.state_dict().items()},
                "step": 0, "lst": [torch.zeros(5, dtype=torch.int64), ""]}
        out = h.load_state_dict(into=into)
        assert out["step"] == 7 and out["lst"][1] == "s"
        assert torch.equal(out["lst"][0], torch.arange(5))
        for k, v in model.state_dict().items():
            assert torch.equal(into["model"][k], v)
        with pytest.raises(ValueError):
            h.load_state_dict(into={"model": {}})
        fresh = h.load_state_dict(device="cpu")
        assert torch.equal(fresh["model"]["fc1.weight"], model.fc1.weight)
        # writing flag set: nothing valid to load
        meta = h.metadata.get()
        meta[DLROVER_CKPT_CONFIG_KEY].writing_shm = True
        h.metadata.set(meta)
        assert h.load_state_dict() == {}
    finally:
        h.unlink()
        h.close()


# ---------------------------------------------------------------- storage
def test_storage_and_deletion(tmp_path):
    st = PosixDiskStorage()
    p = tmp_path / "a.txt"
    st.write("hello", str(p))
    assert st.read(str(p)) == "hello"
    st.write(b"\x00\x01", str(tmp_path / "b.bin"))
    assert st.read(str(tmp_path / "missing")) == ""
    sd = {"x": torch.arange(3)}
    st.write_state_dict(sd, str(tmp_path / "d" / "s.pt"), torch.save)
    back = st.read_state_dict(str(tmp_path / "d" / "s.pt"), lambda q: torch.load(q))
    assert torch.equal(back["x"], sd["x"])
    assert st.get_class_meta().class_name == "PosixDiskStorage"

    for s in (100, 200, 300):
        os.makedirs(tmp_path / str(s))
    tracker = str(tmp_path / CheckpointConstant.TRACER_FILE_NAME)
    keep = PosixStorageWithDeletion(CheckpointConstant.TRACER_FILE_NAME,
                                    KeepStepIntervalStrategy(200, str(tmp_path)))
    keep.write("100", tracker)
    keep.write("200", tracker)
    keep.commit(200, True)  # previous step 100 is not a multiple of 200
    assert not (tmp_path / "100").exists() and (tmp_path / "200").exists()
    latest = KeepLatestStepStrategy(2, str(tmp_path))
    latest.clean_up(200, lambda d: os.rmdir(d))
    latest.clean_up(300, lambda d: os.rmdir(d))
    assert not (tmp_path / "200").exists() and (tmp_path / "300").exists()


# ------------------------------------------------- engine <-> saver (in-proc)
@pytest.fixture
def saver_env():
    os.environ[NodeEnv.NODE_NUM] = "1"
    os.environ[NodeEnv.NODE_RANK] = "0"
    AsyncCheckpointSaver._saver_instance = None
    AsyncCheckpointSaver.start_async_saving_ckpt()
    yield
    if AsyncCheckpointSaver._saver_instance:
        AsyncCheckpointSaver._saver_instance.close()
        AsyncCheckpointSaver._saver_instance = None
    clear_sock_dir()
    os.environ.pop(NodeEnv.NODE_NUM, None)
    os.environ.pop(NodeEnv.NODE_RANK, None)


def _wait_for(pred, timeout=20):
    t0 = time.time()
    while time.time() - t0 < timeout:
        if pred():
            return True
        time.sleep(0.1)
    return False


class SimpleShardingSaver(TempDirCheckpointSaver):
    def persist_to_storage(self, local_shard_id, ckpt_config):
        state_dict = self._shm_handlers[local_shard_id].load_state_dict()
        for name, sd in state_dict.items():
            if name in ckpt_config.paths:
                torch.save(sd, ckpt_config.paths[name])

    def update_tracker_file(self, step):
        with open(os.path.join(self.checkpoint_dir, "tracker.txt"), "w") as f:
            f.write(str(step))


class SimpleShardingCheckpointEngine(FullCheckpointEngine):
    def get_saver_class(self):
        return SimpleShardingSaver

    def get_global_shard_num(self):
        return 1

    def get_local_shard_num(self):
        return 1


def test_engine_save_to_storage_tempdir(saver_env):
    # checkpoint_egine_test.py:147-175
    model = SimpleNet()
    step = 100
    state_dict = dict(model=model.state_dict(), step=step)
    storage = PosixDiskStorage()
    with tempfile.TemporaryDirectory() as tmpdir:
        engine = SimpleShardingCheckpointEngine(tmpdir, storage)
        saved_file = Path(tmpdir) / "checkpoint-100/checkpoint.pt"
        sd = {CheckpointConstant.MODEL_STATES_NAME: state_dict}
        paths = {CheckpointConstant.MODEL_STATES_NAME: saved_file}
        assert engine.save_to_storage(step, sd, paths)
        tracker = Path(tmpdir) / "tracker.txt"
        assert _wait_for(lambda: tracker.exists() and saved_file.exists())
        assert tracker.read_text() == "100"
        assert torch.load(saved_file)["step"] == step
        engine.close()


def test_ddp_checkpointer_memory_and_disk(saver_env):
    # ddp_checkpointer_test.py:58-74 + checkpoint_egine_test.py:225-250
    model = SimpleNet()
    opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.001)
    with tempfile.TemporaryDirectory() as tmpdir:
        ckpt = DdpCheckpointer(tmpdir, deletion_strategy=KeepLatestStepStrategy(1, tmpdir))
        assert isinstance(ckpt.storage, PosixStorageWithDeletion)
        sd = {"model": model.state_dict(), "optimizer": opt.state_dict(), "step": 100}
        assert ckpt.save_checkpoint(100, sd, storage_type=StorageType.MEMORY)
        assert ckpt._engine._shm_handler._buffer_size == 9640
        assert ckpt._engine._shm_handler.shared_memory.size == 9640
        back = ckpt.load_checkpoint()
        assert back["step"] == 100
        for k, v in model.state_dict().items():
            assert torch.equal(back["model"][k], v)
        # in-place variant
        into = {"model": {k: torch.zeros_like(v) for k, v in model.state_dict().items()},
                "optimizer": opt.state_dict(), "step": 0}
        back = ckpt.load_checkpoint(into=into)
        assert torch.equal(into["model"]["fc2.bias"], model.fc2.bias) and back["step"] == 100

        with torch.no_grad():
            model.fc1.weight.add_(1.0)
        sd = {"model": model.state_dict(), "optimizer": opt.state_dict(), "step": 200}
        assert ckpt.save_checkpoint(200, sd, storage_type=StorageType.DISK)
        tracker = os.path.join(tmpdir, CheckpointConstant.TRACER_FILE_NAME)
        assert _wait_for(lambda: os.path.exists(tracker)
                         and open(tracker).read() == "200")
        on_disk = torch.load(os.path.join(tmpdir, "200", "rank_0.pt"))
        assert torch.equal(on_disk["model"]["fc1.weight"], model.fc1.weight)
        with pytest.raises(ValueError):
            ckpt.save_checkpoint(300, sd, storage_type="nope")
        # memory gone -> falls back to storage via the tracker file
        saver = AsyncCheckpointSaver.get_ckpt_saver()
        assert isinstance(saver, DdpCheckpointSaver)
        ckpt._engine._shm_handler.metadata.set({})
        from_disk = ckpt.load_checkpoint()
        assert from_disk["step"] == 200
        ckpt.close()


def test_async_save_holds_lock_until_published(saver_env):
    model = SimpleNet()
    with tempfile.TemporaryDirectory() as tmpdir:
        ckpt = DdpCheckpointer(tmpdir, async_save=True)
        sd = {"model": model.state_dict(), "step": 1}
        assert ckpt.save_checkpoint(1, sd, storage_type=StorageType.MEMORY)
        ckpt.wait_latest_checkpoint()
        assert not ckpt._engine._shm_lock.locked()
        assert ckpt.load_checkpoint()["step"] == 1
        ckpt.close()


def test_saver_breakpoint_save_and_skip_when_locked(saver_env):
    # test_ckpt_saver.py:242-265 (skip logic) and SIGTERM-path persistence
    model = SimpleNet()
    with tempfile.TemporaryDirectory() as tmpdir:
        ckpt = DdpCheckpointer(tmpdir)
        sd = {"model": model.state_dict(), "step": 50}
        ckpt.save_checkpoint(50, sd, storage_type=StorageType.MEMORY)
        saver = AsyncCheckpointSaver.get_ckpt_saver()
        assert _wait_for(lambda: AsyncCheckpointSaver.get_ckpt_saver() is not None)
        saver = AsyncCheckpointSaver.get_ckpt_saver()
        saver._shm_locks[0].acquire()
        saver.save_shm_to_storage()  # a held lock = trainer died mid-copy: skip
        assert not os.path.exists(os.path.join(tmpdir, "50", "rank_0.pt"))
        saver._shm_locks[0].release()
        saver.save_shm_to_storage()
        assert os.path.exists(os.path.join(tmpdir, "50", "rank_0.pt"))
        assert saver._latest_step == 50
        saver._event_queue.put(CheckpointEvent(type=CheckpointEventType.UPDATE_SHARD,
                                               global_shard_num=3))
        assert _wait_for(lambda: saver.global_shard_num == 3)
        ckpt.close()


def test_meta_delta_updates_between_trainer_and_agent():
    """Second and later saves of an unchanged layout send only config,
    checksums and refreshed non-tensor leaves by path; the agent's copy must
    still equal the trainer's, and a layout change falls back to a full set."""
    agent = SharedMemoryHandler(34, host=True)     # serves the meta dict
    trainer = SharedMemoryHandler(34, host=False)  # client, as in a training process
    try:
        sets = []
        real_set = trainer.metadata.set
        trainer.metadata.set = lambda d: (sets.append(1), real_set(d))[1]
        model = SimpleNet()
        opt = torch.optim.SGD(model.parameters(), lr=0.5)
        sd = {"model": model.state_dict(), "optimizer": opt.state_dict(), "step": 1,
              DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=1, paths={"model_states": "/a"})}
        trainer.save_state_dict(sd)
        assert len(sets) == 2  # new layout: full meta, twice (writing, done)
        opt.param_groups[0]["lr"] = 0.25
        sd = {"model": model.state_dict(), "optimizer": opt.state_dict(), "step": 2,
              DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=2, paths={"model_states": "/b"})}
        trainer.save_state_dict(sd)
        assert len(sets) == 2  # nothing but path updates this time
        served = agent.metadata.get()
        assert served["step"] == 2
        assert served[DLROVER_CKPT_CONFIG_KEY].step == 2
        assert served[DLROVER_CKPT_CONFIG_KEY].paths == {"model_states": "/b"}
        assert served[DLROVER_CKPT_CONFIG_KEY].writing_shm is False
        assert served["optimizer"]["param_groups"][0]["lr"] == 0.25
        assert served == trainer.metadata.get(local=True)
        out = agent.load_state_dict()
        assert out["step"] == 2 and torch.equal(out["model"]["fc1.weight"], model.fc1.weight)
        # layout change -> full set again
        sd["extra"] = torch.ones(3)
        sd[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=3, paths={})
        trainer.save_state_dict(sd)
        assert len(sets) == 4
        assert torch.equal(agent.load_state_dict()["extra"], torch.ones(3))
        # agent restarted (empty dict): the update is refused, the full meta follows
        agent.metadata._dict = {}
        sd[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=4, paths={})
        trainer.save_state_dict(sd)
        assert agent.metadata.get()[DLROVER_CKPT_CONFIG_KEY].step == 4
    finally:
        trainer.close()
        agent.unlink()
        agent.close()


def test_shared_memory_is_populated_in_the_background(monkeypatch):
    """A created segment becomes resident from worker threads; wait_populated
    follows the front in address order and close() joins the workers."""
    import dlrover_wuqiong_b200.common.multi_process as mp
    monkeypatch.setattr(mp, "_POPULATE_BLOCK", 1 << 20)
    name = "fcp_test_populate_%d" % os.getpid()
    size = (7 << 20) + 12345
    shm = SharedMemory(name=name, create=True, size=size)
    try:
        shm.wait_populated(3 << 20)
        shm.wait_populated()
        assert shm._populator is None
        assert bytes(shm.buf[:64]) == b"\0" * 64 and shm.buf[size - 1] == 0
        shm.buf[size - 1] = 7
        other = SharedMemory(name=name)  # attaching does not populate
        assert other._populator is None and other.buf[size - 1] == 7
        other.close()
    finally:
        shm.close()
        shm.unlink()
    # closing right after creation must not leave workers touching unmapped memory
    shm = SharedMemory(name=name, create=True, size=size)
    shm.close()
    shm.unlink()
    # FCP_POPULATE_THREADS=0: no background work at all
    monkeypatch.setenv("FCP_POPULATE_THREADS", "0")
    shm = SharedMemory(name=name, create=True, size=4096)
    assert shm._populator is None
    shm.wait_populated()
    shm.close()
    shm.unlink()


def test_shared_dict_update_walks_attributes_and_is_all_or_nothing():
    """SharedDict.update: a path may pass through a plain object (the DCP
    metadata's ``storage_data`` dict is an attribute), a failing path leaves
    everything untouched, and an empty served dict refuses patches."""
    class Box(object):
        def __init__(self):
            self.storage_data = {"a": 1, "b": 2}

    server = SharedDict(name="fcp_attr_%d" % os.getpid(), create=True)
    client = SharedDict(name="fcp_attr_%d" % os.getpid(), create=False)
    try:
        assert client.update([(("x",), 1)]) is False  # nothing served yet
        server._dict = {"meta": Box(), "lst": [10, 20], "cfg": {"step": 1}}
        assert client.update([(("meta", "storage_data", "b"), 22), (("lst", 1), 21),
                              (("cfg", "step"), 2), (("new_key",), "n")]) is True
        d = server._dict
        assert d["meta"].storage_data == {"a": 1, "b": 22} and d["lst"] == [10, 21]
        assert d["cfg"] == {"step": 2} and d["new_key"] == "n"
        # one bad path (list index out of range) -> nothing of the batch is applied
        assert client.update([(("cfg", "step"), 3), (("lst", 5), 0)]) is False
        assert d["cfg"] == {"step": 2}
        # the last container must be a dict or a list: objects are not assigned into
        assert client.update([(("meta", "storage_data"), {})]) is False
        assert d["meta"].storage_data == {"a": 1, "b": 22}
    finally:
        client.unlink()
        server.unlink()


def test_shared_dict_get_keys_fetches_a_subset_without_replacing_the_local_copy():
    server = SharedDict(name="fcp_keys_%d" % os.getpid(), create=True)
    client = SharedDict(name="fcp_keys_%d" % os.getpid(), create=False)
    try:
        client.set({"cfg": {"step": 5}, "big": list(range(1000)), "x": 1})
        local_before = client.get(local=True)
        assert client.get_keys(["cfg", "missing"]) == {"cfg": {"step": 5}}
        assert server.get_keys(["x"]) == {"x": 1}
        assert client.get(local=True) is local_before
        # a server that does not know the request: the whole dict is fetched instead
        real = server._handle

        def old_server(request):
            if request.method == "get_keys":
                raise ValueError("unknown method")
            return real(request)

        server._handle = old_server
        assert client.get_keys(["cfg"]) == {"cfg": {"step": 5}}
    finally:
        client.unlink()
        server.unlink()


def test_reference_named_copy_helpers():
    """_traverse_copy_to_shm / _write_shared_memory (SAVER:174-206) under the
    reference's names: same bytes as the oracle's layout, scalars refreshed."""
    from dlrover_wuqiong_b200 import plan
    from dlrover_wuqiong_b200.ckpt_saver import _traverse_copy_to_shm, _write_shared_memory
    from oracle import oracle
    g = torch.Generator().manual_seed(4)
    sd = {"a": {"w": torch.randn(5, 3, generator=g), "n": 1}, "l": [torch.arange(7), "x"],
          "e": torch.empty(0)}
    meta, leaves, total = plan.build_layout(sd)
    buf = bytearray(total)
    sd["a"]["n"] = 2
    _traverse_copy_to_shm(sd, meta, buf)
    _, _, expect = oracle.pack(sd)
    assert bytes(buf) == expect.tobytes() and meta["a"]["n"] == 2
    buf2 = bytearray(total)
    for lf in leaves:
        _write_shared_memory(lf.tensor, lf.meta, buf2)
    assert buf2 == buf
    with pytest.raises(KeyError):
        _traverse_copy_to_shm({"a": {"w": sd["a"]["w"], "extra": torch.ones(1)}}, meta, buf)
