"""CPU, world_size 2 over gloo: the N>1 plumbing of the path — DDP
checkpointer with rank votes (config 1 of BASELINE.json, scaled down unless
FCP_FULL_CONFIG1=1), FSDP sharded checkpoint through the DCP writer/reader,
and recovery of a replaced rank from a peer.  Pattern of the reference's
mp.spawn tests (dlrover/trainer/tests/torch/checkpoint_egine_test.py:316-327,
checkpoint_backup_test.py:53-123)."""
import os
import socket
import sys
import tempfile
import time
import traceback

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, run_id, fn_name, tmpdir, errq):
    try:
        if ROOT not in sys.path:
            sys.path.insert(0, ROOT)
        os.environ.update({
            "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
            "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank), "LOCAL_WORLD_SIZE": str(world),
            "GROUP_RANK": "0", "GROUP_WORLD_SIZE": "1", "NODE_NUM": "1", "NODE_RANK": "0",
            "ROLE_NAME": "dlrover-trainer", "TORCHELASTIC_RUN_ID": run_id,
            "DLROVER_LOG_LEVEL": "WARNING"})
        torch.manual_seed(1234 + rank)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from dlrover_wuqiong_b200.ckpt_saver import AsyncCheckpointSaver
        if rank == 0:  # local rank 0 plays the agent: hosts the saver thread
            AsyncCheckpointSaver.start_async_saving_ckpt()
        globals()[fn_name](rank, world, tmpdir)
        dist.barrier()
        if rank == 0 and AsyncCheckpointSaver.get_ckpt_saver():
            AsyncCheckpointSaver.get_ckpt_saver().close()
        dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        errq.put("rank %d:\n%s" % (rank, traceback.format_exc()))
        raise


def _spawn(fn_name, world=2):
    ctx = mp.get_context("spawn")
    errq = ctx.SimpleQueue()
    port, run_id = _free_port(), "fcp_dist_%d_%s" % (os.getpid(), fn_name)
    with tempfile.TemporaryDirectory() as tmpdir:
        procs = [ctx.Process(target=_worker, args=(r, world, port, run_id, fn_name, tmpdir, errq))
                 for r in range(world)]
        for p in procs:
            p.start()
        deadline = time.time() + 240
        for p in procs:
            p.join(max(1, deadline - time.time()))
        alive = [p for p in procs if p.is_alive()]
        for p in alive:
            p.kill()
        errors = []
        while not errq.empty():
            errors.append(errq.get())
        assert not alive, "workers hung"
        assert not errors and all(p.exitcode == 0 for p in procs), "\n".join(errors)


def _wait_for(pred, timeout=30):
    t0 = time.time()
    while time.time() - t0 < timeout:
        if pred():
            return True
        time.sleep(0.2)
    return False


# ------------------------------------------------------------------ models
class Block(nn.Module):
    def __init__(self, e):
        super().__init__()
        self.ln_1 = nn.LayerNorm(e, bias=False)
        self.attn = nn.Linear(e, 3 * e, bias=False)
        self.proj = nn.Linear(3 * e, e, bias=False)
        self.ln_2 = nn.LayerNorm(e, bias=False)
        self.fc = nn.Linear(e, 4 * e, bias=False)
        self.fc2 = nn.Linear(4 * e, e, bias=False)

    def forward(self, x):
        x = x + self.proj(self.attn(self.ln_1(x)))
        return x + self.fc2(torch.relu(self.fc(self.ln_2(x))))


class TinyGPT(nn.Module):
    """nanoGPT-shaped (tied wte / lm_head, bias=False)."""

    def __init__(self, n_layer, e, vocab, block):
        super().__init__()
        self.wte = nn.Embedding(vocab, e)
        self.wpe = nn.Embedding(block, e)
        self.h = nn.ModuleList([Block(e) for _ in range(n_layer)])
        self.ln_f = nn.LayerNorm(e, bias=False)
        self.lm_head = nn.Linear(e, vocab, bias=False)
        self.wte.weight = self.lm_head.weight

    def forward(self, idx):
        pos = torch.arange(idx.shape[1])
        x = self.wte(idx) + self.wpe(pos)
        for b in self.h:
            x = b(x)
        return self.lm_head(self.ln_f(x))


def _gpt_cfg():
    if os.getenv("FCP_FULL_CONFIG1") == "1":
        return dict(n_layer=12, e=768, vocab=50304, block=128)  # GPT-2 small 124M
    return dict(n_layer=2, e=64, vocab=503, block=16)


# ------------------------------------------------------------------- cases
def ddp_case(rank, world, tmpdir):
    """config 1: DdpCheckpointer save/load to shared memory, ws=2, gloo."""
    from dlrover_wuqiong_b200.common.constants import CheckpointConstant
    from dlrover_wuqiong_b200.flash_checkpoint.ddp import DdpCheckpointer, StorageType
    from oracle import oracle
    torch.manual_seed(7)  # replicas hold identical weights
    cfg = _gpt_cfg()
    model = TinyGPT(**cfg)
    ddp = nn.parallel.DistributedDataParallel(model)
    opt = torch.optim.AdamW(ddp.parameters(), lr=1e-3)
    ddp(torch.randint(0, cfg["vocab"], (2, cfg["block"]))).sum().backward()
    opt.step()
    ckpt = DdpCheckpointer(tmpdir)  # replicated: only local rank 0 writes
    sd = {"model": model.state_dict(), "optimizer": opt.state_dict(), "step": 10}
    saved = ckpt.save_checkpoint(10, sd, storage_type=StorageType.MEMORY)
    assert saved == (rank == 0)
    if rank == 0:
        _, total, expect = oracle.pack({CheckpointConstant.MODEL_STATES_NAME: sd})
        shm = ckpt._engine._shm_handler.shared_memory
        assert shm.size == total and bytes(shm.buf[:total]) == expect.tobytes()
    dist.barrier()
    back = ckpt.load_checkpoint()  # every rank reads shard 0's segment
    assert back["step"] == 10
    for k, v in model.state_dict().items():
        assert torch.equal(back["model"][k], v), k
    assert torch.equal(back["optimizer"]["state"][0]["exp_avg"],
                       opt.state_dict()["state"][0]["exp_avg"])
    # DISK: one SAVE event from global rank 0; node 0 persists rank_0.pt
    ckpt.save_checkpoint(20, sd, storage_type=StorageType.DISK)
    if rank == 0:
        tracker = os.path.join(tmpdir, CheckpointConstant.TRACER_FILE_NAME)
        assert _wait_for(lambda: os.path.exists(tracker) and open(tracker).read() == "20")
        on_disk = torch.load(os.path.join(tmpdir, "20", "rank_0.pt"), weights_only=False)
        assert torch.equal(on_disk["model"]["wpe.weight"], model.state_dict()["wpe.weight"])
    dist.barrier()
    ckpt.close()


def fsdp_case(rank, world, tmpdir):
    """FSDP sharded state (model + AdamW) through the DCP shm writer/reader."""
    from torch.distributed.fsdp import FullyShardedDataParallel as FSDP
    from dlrover_wuqiong_b200.ckpt_saver import DLROVER_CKPT_CONFIG_KEY
    from dlrover_wuqiong_b200.common.constants import CheckpointConstant
    from dlrover_wuqiong_b200.flash_checkpoint.fsdp import FsdpCheckpointer, StorageType
    torch.manual_seed(11)
    cfg = dict(n_layer=2, e=32, vocab=101, block=8)
    model = FSDP(TinyGPT(**cfg), device_id=torch.device("cpu"))
    opt = torch.optim.AdamW(model.parameters(), lr=1e-2)
    for _ in range(2):
        opt.zero_grad()
        model(torch.randint(0, cfg["vocab"], (2, cfg["block"]))).sum().backward()
        opt.step()
    flat_before = [p.detach().clone() for p in model.parameters()]
    opt_before = {i: {k: (v.clone() if torch.is_tensor(v) else v) for k, v in s.items()}
                  for i, s in enumerate(opt.state.values())}
    ckpt = FsdpCheckpointer(tmpdir)
    assert ckpt.save_checkpoint(5, model, opt, {"step": 5, "epoch": 3},
                                storage_type=StorageType.MEMORY)
    handler = ckpt._engine._shm_handler
    meta = handler.metadata.get()
    conf = meta[DLROVER_CKPT_CONFIG_KEY]
    assert conf.step == 5 and not conf.writing_shm
    assert conf.paths[CheckpointConstant.MODEL_STATES_NAME].endswith("__%d_0.distcp" % rank)
    # dense layout: this rank's items tile its segment exactly
    mine = sorted((i for i in meta["dcp_metadata"].storage_data.values()
                   if i.relative_path == "__%d_0.distcp" % rank), key=lambda i: i.offset)
    assert mine and mine[0].offset == 0
    for a, b in zip(mine, mine[1:]):
        assert a.offset + a.length == b.offset
    assert mine[-1].offset + mine[-1].length == handler.shared_memory.size
    # wreck the live state, then restore from memory
    with torch.no_grad():
        for p in model.parameters():
            p.zero_()
    for s in opt.state.values():
        for v in s.values():
            if torch.is_tensor(v):
                v.zero_()
    extra = ckpt.load_checkpoint(model, opt)
    assert extra["step"] == 5  # only the keys the loader asks for come back (fsdp.py:120-125)
    for a, b in zip(flat_before, model.parameters()):
        assert torch.equal(a, b.detach())
    for i, s in enumerate(opt.state.values()):
        for k, v in s.items():
            if torch.is_tensor(v):
                assert torch.equal(v, opt_before[i][k]), (i, k)
    # DISK: every rank's segment is written raw, node 0 adds .metadata + tracker
    ckpt.save_checkpoint(6, model, opt, {"step": 6}, storage_type=StorageType.DISK)
    assert ckpt._engine.last_save_planned is True  # "epoch" is gone: the local plans changed
    if rank == 0:
        d = os.path.join(tmpdir, "6")
        assert _wait_for(lambda: os.path.isdir(d) and sorted(os.listdir(d)) ==
                         [".metadata", "__0_0.distcp", "__1_0.distcp"])
        tracker = os.path.join(tmpdir, CheckpointConstant.TRACER_FILE_NAME)
        assert _wait_for(lambda: os.path.exists(tracker) and open(tracker).read() == "6")
    dist.barrier()
    # storage fallback: FileReader over the persisted files
    from dlrover_wuqiong_b200.flash_checkpoint.fsdp_engine import FileReader
    handler.metadata.set({})
    dist.barrier()
    with torch.no_grad():
        for p in model.parameters():
            p.zero_()
    reader = ckpt._engine.load()
    assert isinstance(reader, FileReader)
    extra = ckpt.load_checkpoint(model, opt)
    assert extra["step"] == 6
    for a, b in zip(flat_before, model.parameters()):
        assert torch.equal(a, b.detach())
    # cached plan with new values: train on, save (the pickled `step` grows from
    # 1 to 3 digits+, so rank 0's item offsets move), wreck, restore from memory
    opt.zero_grad()
    model(torch.randint(0, cfg["vocab"], (2, cfg["block"]))).sum().backward()
    opt.step()
    flat_now = [p.detach().clone() for p in model.parameters()]
    opt_now = {i: {k: (v.clone() if torch.is_tensor(v) else v) for k, v in s.items()}
               for i, s in enumerate(opt.state.values())}
    assert any(not torch.equal(a, b) for a, b in zip(flat_before, flat_now))
    assert ckpt.save_checkpoint(70000, model, opt, {"step": 70000},
                                storage_type=StorageType.MEMORY)
    assert ckpt._engine.last_save_planned is False
    meta = handler.metadata.get()
    mine = sorted((i for i in meta["dcp_metadata"].storage_data.values()
                   if i.relative_path == "__%d_0.distcp" % rank), key=lambda i: i.offset)
    for a, b in zip(mine, mine[1:]):
        assert a.offset + a.length == b.offset
    assert mine[-1].offset + mine[-1].length == handler.shared_memory.size
    with torch.no_grad():
        for p in model.parameters():
            p.zero_()
    for s in opt.state.values():
        for v in s.values():
            if torch.is_tensor(v):
                v.zero_()
    extra = ckpt.load_checkpoint(model, opt)
    assert extra["step"] == 70000
    for a, b in zip(flat_now, model.parameters()):
        assert torch.equal(a, b.detach())
    for i, s in enumerate(opt.state.values()):
        for k, v in s.items():
            if torch.is_tensor(v):
                assert torch.equal(v, opt_now[i][k]), (i, k)
    # a new key changes the local plans: planned again, and restorable
    assert ckpt.save_checkpoint(70001, model, opt, {"step": 70001, "epoch": 4},
                                storage_type=StorageType.MEMORY)
    assert ckpt._engine.last_save_planned is True
    with torch.no_grad():
        for p in model.parameters():
            p.zero_()
    extra = ckpt.load_checkpoint(model, opt)
    assert extra["step"] == 70001
    for a, b in zip(flat_now, model.parameters()):
        assert torch.equal(a, b.detach())
    # cached plan + patched meta all the way to storage: what the agent persists
    # must be a loadable DCP checkpoint of the latest values
    ckpt.save_checkpoint(70002, model, opt, {"step": 70002, "epoch": 5},
                         storage_type=StorageType.DISK)
    assert ckpt._engine.last_save_planned is False
    if rank == 0:
        d = os.path.join(tmpdir, "70002")
        assert _wait_for(lambda: os.path.isdir(d) and sorted(os.listdir(d)) ==
                         [".metadata", "__0_0.distcp", "__1_0.distcp"])
        tracker = os.path.join(tmpdir, CheckpointConstant.TRACER_FILE_NAME)
        assert _wait_for(lambda: os.path.exists(tracker) and open(tracker).read() == "70002")
    dist.barrier()
    handler.metadata.set({})
    dist.barrier()
    with torch.no_grad():
        for p in model.parameters():
            p.zero_()
    assert isinstance(ckpt._engine.load(), FileReader)
    extra = ckpt.load_checkpoint(model, opt)
    assert extra["step"] == 70002, extra
    for a, b in zip(flat_now, model.parameters()):
        assert torch.equal(a, b.detach())
    ckpt.close()


def peer_recovery_case(rank, world, tmpdir):
    """A replaced rank (rank 1, nothing in memory) gets the checkpoint image
    from the surviving replica (rank 0) — host transport (gather) and the
    device-transport code path run over gloo with CPU tensors."""
    from dlrover_wuqiong_b200.ckpt_saver import (DLROVER_CKPT_CONFIG_KEY, FCP_CHECKSUM_KEY,
                                                 CheckpointConfig, SharedMemoryHandler)
    from dlrover_wuqiong_b200.flash_checkpoint.replica import (FullCkptReplicaManager,
                                                               ShardCkptReplicaManager)
    torch.manual_seed(5)
    model = nn.Sequential(nn.Linear(16, 16), nn.ReLU(), nn.Linear(16, 8))
    opt = torch.optim.Adam(model.parameters(), lr=0.1)
    # each "node" has one rank here
    os.environ.update({"NODE_NUM": "2", "NODE_RANK": str(rank), "LOCAL_WORLD_SIZE": "1",
                       "LOCAL_RANK": "0"})
    handler = SharedMemoryHandler(10 + rank, host=True)
    sd = {"model": model.state_dict(), "optimizer": opt.state_dict(),
          DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(rank=rank, step=7, world_size=world)}
    if rank == 0:
        handler.save_state_dict(sd)
        assert handler.shared_memory.size == 1632  # checkpoint_backup_test.py:86
    mgr = FullCkptReplicaManager(replica_count=1)
    assert mgr.backup_ranks == [0, 1]
    image, meta = mgr.gather(handler)
    assert image.numel() == 1632 and meta[DLROVER_CKPT_CONFIG_KEY].step == 7
    if rank == 0:
        assert bytes(handler.shared_memory.buf[:1632]) == image.numpy().tobytes()
    # replica_count == 0: inert, and no ZeroDivisionError for sharded engines
    assert ShardCkptReplicaManager(0).backup_ranks == []
    assert ShardCkptReplicaManager(0).gather(handler) == (None, None)
    # sharded backup group: each rank ends up holding its peer's shard too
    h2 = SharedMemoryHandler(20 + rank, host=True)
    mine = {"w": torch.full((4, 4), float(rank + 1)),
            DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(rank=rank, step=9, world_size=world)}
    h2.save_state_dict(mine)
    smgr = ShardCkptReplicaManager(replica_count=2)
    assert smgr.backup_ranks == [0, 1]
    smgr.backup(h2)
    peer = 1 - rank
    peer_handler = smgr._rank_shms[peer]
    got = peer_handler.load_state_dict()
    assert torch.equal(got["w"], torch.full((4, 4), float(peer + 1)))
    for h in (handler, h2, peer_handler):
        h.unlink()
        h.close()


@pytest.mark.timeout(300)
def test_ddp_checkpointer_two_ranks_gloo():
    _spawn("ddp_case")


@pytest.mark.timeout(300)
def test_fsdp_sharded_checkpoint_two_ranks_gloo():
    _spawn("fsdp_case")


@pytest.mark.timeout(300)
def test_peer_recovery_two_ranks_gloo():
    _spawn("peer_recovery_case")
