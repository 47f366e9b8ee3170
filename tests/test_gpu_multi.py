"""GPU, 2 ranks over NCCL (skipped on a 1-GPU box): recovery of a replaced
rank of a replicated DDP group from the packed image a surviving peer still
holds in HBM — NCCL broadcast over NVLink, then the scatter kernel with CRC
verification (replaces the gloo/CPU broadcast of
dlrover/trainer/torch/flash_checkpoint/replica.py:275-332) — and every rank
checkpointing its own shard concurrently."""
import os
import socket
import sys
import tempfile
import traceback

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, run_id, tmpdir, errq):
    try:
        if ROOT not in sys.path:
            sys.path.insert(0, ROOT)
        os.environ.update({
            "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
            "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank), "LOCAL_WORLD_SIZE": str(world),
            "NODE_NUM": "1", "NODE_RANK": "0", "ROLE_NAME": "dlrover-trainer",
            "TORCHELASTIC_RUN_ID": run_id, "DLROVER_LOG_LEVEL": "WARNING"})
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world,
                                device_id=torch.device("cuda", rank))
        from dlrover_wuqiong_b200.ckpt_saver import (DLROVER_CKPT_CONFIG_KEY,
                                                     AsyncCheckpointSaver, CheckpointConfig,
                                                     SharedMemoryHandler)
        from dlrover_wuqiong_b200.flash_checkpoint.ddp import DdpCheckpointer, StorageType
        from dlrover_wuqiong_b200.flash_checkpoint.replica import FullCkptReplicaManager
        if rank == 0:
            AsyncCheckpointSaver.start_async_saving_ckpt()

        # --- every rank saves / restores its own shard at the same time
        torch.manual_seed(100 + rank)
        model = nn.Sequential(nn.Linear(1024, 2048), nn.Linear(2048, 512)).cuda()
        opt = torch.optim.AdamW(model.parameters(), lr=1e-3, fused=True)
        model(torch.randn(4, 1024, device="cuda")).sum().backward()
        opt.step()
        ckpt = DdpCheckpointer(tmpdir, local_shard_num=world, global_shard_num=world)
        sd = {"model": model.state_dict(), "optimizer": opt.state_dict(), "step": 3}
        want = {k: v.clone() for k, v in model.state_dict().items()}
        assert ckpt.save_checkpoint(3, sd, storage_type=StorageType.MEMORY)
        with torch.no_grad():
            for p in model.parameters():
                p.zero_()
        ckpt.load_checkpoint(into=sd)
        for k, v in want.items():
            assert torch.equal(model.state_dict()[k], v), k
        ckpt.close()

        # --- replicated group: rank 1 was replaced and has nothing in memory
        torch.manual_seed(7)
        rep = nn.Sequential(nn.Linear(777, 1333), nn.Linear(1333, 91)).cuda()
        ropt = torch.optim.AdamW(rep.parameters(), lr=1e-3, fused=True)
        rep(torch.randn(4, 777, device="cuda")).sum().backward()
        ropt.step()
        rsd = {"model": rep.state_dict(), "optimizer": ropt.state_dict(), "step": 11}
        rwant = {k: v.clone() for k, v in rep.state_dict().items()}
        handler = SharedMemoryHandler(10 + rank, host=True)
        if rank == 0:
            full = dict(rsd)
            full[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=11, rank=0, world_size=world)
            handler.save_state_dict(full)
        else:
            with torch.no_grad():  # the replacement starts from garbage
                for p in rep.parameters():
                    p.fill_(-1.0)
        mgr = FullCkptReplicaManager(replica_count=0)
        step = mgr.recover_on_device(handler, rsd, healthy=(rank == 0))
        assert step == 11
        for k, v in rwant.items():
            assert torch.equal(rep.state_dict()[k], v), k
        # --- same recovery, fused: the replaced rank's scatter kernel reads the
        # survivor's staging image over NVLink (CUDA IPC), no receive buffer
        if rank == 1:
            with torch.no_grad():
                for p in rep.parameters():
                    p.fill_(-2.0)
        step = mgr.recover_fused(handler, rsd, healthy=(rank == 0))
        assert step == 11
        for k, v in rwant.items():
            assert torch.equal(rep.state_dict()[k], v), k

        # --- rates on a 2 GiB replicated state (recorded, not asserted)
        import json
        import time
        big = {"w%d" % i: torch.randn(64 << 20, device="cuda") for i in range(8)}  # 2 GiB
        bwant = big["w3"][:1024].clone() if rank == 0 else None
        bh = SharedMemoryHandler(20 + rank, host=True)
        if rank == 0:
            full = dict(big)
            full[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=5, rank=0, world_size=world)
            bh.save_state_dict(full)
        rates = {}
        for name, fn in (("nccl_broadcast_then_scatter", mgr.recover_on_device),
                         ("fused_p2p_scatter", mgr.recover_fused)):
            for it in range(3):
                if rank == 1:
                    big["w3"].zero_()
                torch.cuda.synchronize()
                dist.barrier()
                t0 = time.perf_counter()
                assert fn(bh, big, healthy=(rank == 0)) == 5
                torch.cuda.synchronize()
                dt = time.perf_counter() - t0
                rates[name] = min(rates.get(name, 1e9), dt)
            ref = torch.empty(1024, device="cuda")
            if rank == 0:
                ref.copy_(bwant)
            dist.broadcast(ref, src=0)
            assert torch.equal(big["w3"][:1024], ref)
        if rank == 1:
            total = 8 * (64 << 20) * 4
            out = {k: {"s": v, "GBps": total / v / 1e9} for k, v in rates.items()}
            out["bytes"] = total
            out["fused_breakdown"] = getattr(mgr, "last_fused_stats", None)
            os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
            with open(os.path.join(ROOT, "gpurun_out", "peer_recovery.json"), "w") as f:
                json.dump(out, f)
        dist.barrier()
        if rank == 0:
            bh.unlink()
        bh.close()
        dist.barrier()
        if rank == 0:
            handler.unlink()
            saver = AsyncCheckpointSaver.get_ckpt_saver()
            if saver:
                saver.close()
        handler.close()
        dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        errq.put("rank %d:\n%s" % (rank, traceback.format_exc()))
        raise


def _fsdp_worker(rank, world, port, run_id, tmpdir, errq):
    """FSDP sharded checkpoint on GPUs: planned save, cached-plan saves (sync
    and async), restore through the scatter kernel; rank 0 records how long
    save_to_memory blocks."""
    try:
        import json
        import time
        if ROOT not in sys.path:
            sys.path.insert(0, ROOT)
        os.environ.update({
            "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
            "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank), "LOCAL_WORLD_SIZE": str(world),
            "NODE_NUM": "1", "NODE_RANK": "0", "ROLE_NAME": "dlrover-trainer",
            "TORCHELASTIC_RUN_ID": run_id, "DLROVER_LOG_LEVEL": "WARNING"})
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world,
                                device_id=torch.device("cuda", rank))
        from torch.distributed.fsdp import FullyShardedDataParallel as FSDP
        from dlrover_wuqiong_b200.ckpt_saver import AsyncCheckpointSaver
        from dlrover_wuqiong_b200.flash_checkpoint.fsdp import FsdpCheckpointer, StorageType
        if rank == 0:
            AsyncCheckpointSaver.start_async_saving_ckpt()
        torch.manual_seed(9)
        width, layers = 1024, 24
        blocks = []
        for _ in range(layers):
            blocks += [nn.LayerNorm(width), nn.Linear(width, 4 * width), nn.GELU(),
                       nn.Linear(4 * width, width)]
        model = FSDP(nn.Sequential(*blocks).cuda(), device_id=torch.device("cuda", rank))
        opt = torch.optim.AdamW(model.parameters(), lr=1e-4)

        def train_step():
            opt.zero_grad()
            model(torch.randn(2, width, device="cuda")).square().mean().backward()
            opt.step()

        def snapshot():
            return ([p.detach().clone() for p in model.parameters()],
                    {i: {k: v.clone() for k, v in st.items() if torch.is_tensor(v)}
                     for i, st in enumerate(opt.state.values())})

        def wreck():
            with torch.no_grad():
                for p in model.parameters():
                    p.zero_()
            for st in opt.state.values():
                for v in st.values():
                    if torch.is_tensor(v):
                        v.zero_()

        def check(snap):
            for a, b in zip(snap[0], model.parameters()):
                assert torch.equal(a, b.detach())
            for i, st in enumerate(opt.state.values()):
                for k, v in st.items():
                    if torch.is_tensor(v):
                        assert torch.equal(v, snap[1][i][k]), (i, k)

        timings = []

        def timed_save(ckpt, step):
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            assert ckpt.save_checkpoint(step, model, opt, {"step": step},
                                        storage_type=StorageType.MEMORY)
            dt = time.perf_counter() - t0
            eng = ckpt._engine
            timings.append({"step": step, "s": dt, "planned": eng.last_save_planned,
                            "async": eng._async_save,
                            "save_to_memory_s": getattr(eng, "last_save_to_memory_s", None)})

        train_step()
        for async_save in (False, True):
            ckpt = FsdpCheckpointer(os.path.join(tmpdir, "a%d" % async_save), async_save=async_save)
            base = 100 if async_save else 0
            timed_save(ckpt, base + 1)
            assert ckpt._engine.last_save_planned is True
            for k in (2, 3, 4):
                train_step()
                snap = snapshot()
                timed_save(ckpt, base + k)
                assert ckpt._engine.last_save_planned is False
                if async_save:
                    train_step()  # the live tensors move on while the DMA runs
                    ckpt.wait_latest_checkpoint()
            shard_bytes = ckpt._engine._shm_handler.shared_memory.size
            wreck()
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            extra = ckpt.load_checkpoint(model, opt)
            torch.cuda.synchronize()
            timings.append({"load_checkpoint_s": time.perf_counter() - t0, "async": async_save,
                            "restore_path": ckpt._engine._shm_handler.last_stats.get("restore_path")})
            assert extra["step"] == base + 4
            check(snap)
            ckpt.close()
        if rank == 0:
            os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
            with open(os.path.join(ROOT, "gpurun_out", "fsdp_blocking.json"), "w") as f:
                json.dump({"world": world, "shard_bytes_rank0": shard_bytes,
                           "what": "FsdpCheckpointer.save_checkpoint(MEMORY) wall time incl. "
                                   "FSDP state-dict assembly", "saves": timings}, f, indent=1)
        dist.barrier()
        if rank == 0:
            saver = AsyncCheckpointSaver.get_ckpt_saver()
            if saver:
                saver.close()
        dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        errq.put("rank %d:\n%s" % (rank, traceback.format_exc()))
        raise


def _run_two_ranks(worker, tag, world=2):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs (run with gpurun --gpus %d)" % (world, world))
    ctx = mp.get_context("spawn")
    errq = ctx.SimpleQueue()
    port, run_id = _free_port(), "fcp_gpu%d_%s_%d" % (world, tag, os.getpid())
    with tempfile.TemporaryDirectory() as tmpdir:
        procs = [ctx.Process(target=worker, args=(r, world, port, run_id, tmpdir, errq))
                 for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(200)
        alive = [p for p in procs if p.is_alive()]
        for p in alive:
            p.kill()
        errors = []
        while not errq.empty():
            errors.append(errq.get())
        assert not alive, "workers hung"
        assert not errors and all(p.exitcode == 0 for p in procs), "\n".join(errors)


@pytest.mark.timeout(300)
def test_two_rank_fsdp_sharded_cached_plan_and_async():
    _run_two_ranks(_fsdp_worker, "fsdp")


@pytest.mark.timeout(300)
def test_one_rank_fsdp_checkpointer_cached_plan_and_async():
    """The same through FsdpCheckpointer with a world of one rank (NCCL), so
    that a 1-GPU box also exercises the FSDP engine on CUDA: planned save,
    cached-plan saves (sync and async), restore through the scatter kernel."""
    _run_two_ranks(_fsdp_worker, "fsdp1", world=1)


@pytest.mark.timeout(300)
def test_two_rank_nccl_save_restore_and_peer_recovery():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    ctx = mp.get_context("spawn")
    errq = ctx.SimpleQueue()
    port, run_id = _free_port(), "fcp_gpu2_%d" % os.getpid()
    with tempfile.TemporaryDirectory() as tmpdir:
        procs = [ctx.Process(target=_worker, args=(r, 2, port, run_id, tmpdir, errq))
                 for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(150)
        alive = [p for p in procs if p.is_alive()]
        for p in alive:
            p.kill()
        errors = []
        while not errq.empty():
            errors.append(errq.get())
        assert not alive, "workers hung"
        assert not errors and all(p.exitcode == 0 for p in procs), "\n".join(errors)
