"""FSDP / DCP shm writer and reader against the REFERENCE's writer
(dlrover/trainer/torch/flash_checkpoint/fsdp_engine.py:85-221,239-297): the
fixture tests/golden/ref_fsdp_writer.pt was produced by the reference itself
(oracle/gen_golden_fsdp.py).  Parity is defined at the
``write_data(plan, planner)`` boundary: same file name, same
``_StorageInfo(relative_path, offset, length)`` for every plan item in plan
order, identical bytes in the used prefix of the segment, same non-sharded
data in the meta; and a checkpoint written by the reference (its bytes + its
meta) restores through this package's reader.

CPU variant here; the ``gpu`` variants run the same on CUDA tensors (one pack
launch set / the scatter kernel), world size 1.
"""
import io
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.distributed.checkpoint as dist_cp

os.environ.setdefault("ROLE_NAME", "dlrover-trainer")
os.environ.setdefault("TORCHELASTIC_RUN_ID", "fcp_test_%d" % os.getpid())

from dlrover_wuqiong_b200.ckpt_saver import FCP_CHECKSUM_KEY, SharedMemoryHandler  # noqa: E402
from dlrover_wuqiong_b200.flash_checkpoint import fsdp_engine  # noqa: E402
from oracle import fsdp_case, oracle  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def fixture():
    return torch.load(os.path.join(ROOT, "tests", "golden", "ref_fsdp_writer.pt"),
                      weights_only=False)


@pytest.fixture
def one_rank_group():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    dist.init_process_group("gloo", rank=0, world_size=1,
                            init_method="tcp://127.0.0.1:%d" % port)
    yield
    dist.destroy_process_group()


def _save_with_this_package(fixture, dev, shard_id):
    sd = fsdp_case.build(fixture["inputs"], dev)
    handler = SharedMemoryHandler(shard_id, host=True)
    writer = fsdp_engine.SharedMemoryWriter(handler)
    dist_cp.save(sd, storage_writer=writer,
                 planner=dist_cp.DefaultSavePlanner(dedup_save_to_lowest_rank=True))
    return sd, handler, writer


def _check_against_reference(fixture, handler, writer):
    assert writer.file_name == fixture["file_name"] == "__0_0.distcp"
    md = writer.metadata["dcp_metadata"]
    mine = [(i.fqn, None if i.offset is None else tuple(i.offset), i.index,
             v.relative_path, v.offset, v.length) for i, v in md.storage_data.items()]
    ref = [(it["fqn"], it["offset"], it["index"], it["relative_path"], it["at"], it["length"])
           for it in fixture["items"]]
    same_torch = torch.__version__ == fixture["torch"]
    if same_torch:
        assert mine == ref  # plan order, places and lengths, item for item
    else:  # pickled BYTE_IO lengths move between torch releases: tensors only
        assert [m[:3] for m in mine] == [r[:3] for r in ref]
    used = max(v.offset + v.length for v in md.storage_data.values())
    # this package sizes the segment to what is written (the reference sizes
    # it by the global shapes: same thing on one rank)
    assert handler.shared_memory.size == used
    got = bytes(handler.shared_memory.buf[:used])
    if same_torch:
        assert used == fixture["used"] == fixture["shm_size"]
        assert got == fixture["prefix"]
    else:
        for m, r in zip(mine, ref):
            if r[1] is not None or r[0].endswith((".b", ".step")):
                assert got[m[4]:m[4] + m[5]] == fixture["prefix"][r[4]:r[4] + r[5]], m[0]
    # non-sharded data: same keys; tensors equal (they travel re-encoded as
    # (tag, dtype, shape, bytes) tuples), pickles equal
    mine_ns = writer.metadata["no_shard_data"]
    assert list(mine_ns) == list(fixture["no_shard"])
    for fqn, (kind, value) in fixture["no_shard"].items():
        mine_v = fsdp_engine._expand(fsdp_engine._compact(mine_ns[fqn]))
        if kind == "tensor":
            assert torch.equal(mine_v.cpu(), value), fqn
        elif same_torch:
            assert mine_v.getvalue() == value, fqn
    return used, got


def _reference_checkpoint_loads(fixture, dev, shard_id):
    """The reference's bytes + the reference's meta, read by this package."""
    handler = SharedMemoryHandler(shard_id, host=True)
    handler.init_shared_memory(create=True, size=fixture["shm_size"])
    handler.shared_memory.buf[:fixture["used"]] = fixture["prefix"]
    md = fixture["dcp_metadata"]
    md = dist_cp.metadata.Metadata(
        state_dict_metadata=md.state_dict_metadata, planner_data=md.planner_data,
        storage_data={i: fsdp_engine._StorageInfo(*v) for i, v in md.storage_data.items()})
    no_shard = {k: (io.BytesIO(v) if kind == "bytes" else v)
                for k, (kind, v) in fixture["no_shard"].items()}
    from dlrover_wuqiong_b200.ckpt_saver import DLROVER_CKPT_CONFIG_KEY, CheckpointConfig
    handler.metadata.set({DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=3),
                          "dcp_metadata": md, "no_shard_data": no_shard})
    target = fsdp_case.build(fixture["inputs"], dev, zeros=True)
    target["step"] = 0
    reader = fsdp_engine.SharedMemoryReader(handler)
    dist_cp.load(target, storage_reader=reader, planner=dist_cp.DefaultLoadPlanner())
    want = fsdp_case.local_tensors(fsdp_case.build(fixture["inputs"], "cpu"))
    for name, t in fsdp_case.local_tensors(target).items():
        assert torch.equal(t.cpu(), want[name]), name
    assert target["step"] == 7
    assert target["optim"]["param_groups"][0]["betas"] == (0.9, 0.95)
    return handler


def test_writer_matches_reference_on_host_tensors(fixture, one_rank_group):
    sd, handler, writer = _save_with_this_package(fixture, "cpu", 62)
    try:
        _check_against_reference(fixture, handler, writer)
    finally:
        handler.close()
        handler.unlink()


def test_reader_loads_a_reference_written_checkpoint_on_host(fixture, one_rank_group):
    handler = _reference_checkpoint_loads(fixture, "cpu", 63)
    handler.close()
    handler.unlink()


@pytest.mark.gpu
def test_writer_matches_reference_on_cuda_tensors(fixture, one_rank_group):
    sd, handler, writer = _save_with_this_package(fixture, "cuda:0", 64)
    try:
        used, got = _check_against_reference(fixture, handler, writer)
        # the pack kernel's per-chunk CRC-32C of that image
        entry = writer.checksums
        assert entry["sums"] == oracle.chunk_crc32c(
            memoryview(got), entry["chunk_bytes"]).tolist()
        assert handler.last_stats.get("save_mode") is not None  # went through libfcp
    finally:
        handler.close()
        handler.unlink()


@pytest.mark.gpu
def test_reader_loads_a_reference_written_checkpoint_on_cuda(fixture, one_rank_group):
    handler = _reference_checkpoint_loads(fixture, "cuda:0", 65)
    assert handler.last_stats["restore_path"] in ("registered", "bounced")  # scatter kernel
    handler.close()
    handler.unlink()
