"""GPU, 2 ranks over NCCL (skipped on a 1-GPU box): replicated (DDP) state
through ``DdpCheckpointer`` — the path the reference serves with local rank 0
writing the whole image and every rank reading it back through the host
(full_ckpt_engine.py:75-88,145-166), plus gloo for a lost replica
(replica.py:275-332).

* striped save: both ranks write half of ONE segment; its bytes and chunk sums
  equal the oracle's single-writer image;
* ``load_checkpoint(into=)`` after a rank lost its tensors: stripes come from
  the peers' HBM over NVLink, nothing crosses the host link;
* the same after a restart (nothing in HBM): every rank fetches its half of
  the shm image and the halves are exchanged over NVLink;
* single-writer save, then recovery of the other rank from the survivor's HBM;
* a corrupted stripe is caught by the CRC check on every rank.
"""
import os
import socket
import sys
import tempfile
import traceback

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _state(dev):
    """The same on every rank (a DDP replica)."""
    g = torch.Generator(device=dev)
    g.manual_seed(2024)
    model, opt = {}, {}
    for i, n in enumerate((1 << 20, 4099, 3 << 20, 77, 5 << 20, 1 << 18)):
        w = torch.empty(n, device=dev).normal_(0, 1, generator=g)
        model["w%d" % i] = w if i % 2 else w.to(torch.bfloat16)
        opt[i] = {"step": torch.tensor(float(i), device=dev),
                  "exp_avg": torch.empty(n, device=dev).normal_(0, 1, generator=g),
                  "exp_avg_sq": torch.empty(n, device=dev).uniform_(0, 1, generator=g)}
    return {"model": model, "optimizer": {"state": opt}, "step": 9}


def _tensors(sd):
    out = []

    def walk(v):
        if isinstance(v, dict):
            for x in v.values():
                walk(x)
        elif torch.is_tensor(v):
            out.append(v)

    walk(sd)
    return out


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
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
        import numpy as np
        from dlrover_wuqiong_b200.ckpt_saver import FCP_CHECKSUM_KEY, AsyncCheckpointSaver
        from dlrover_wuqiong_b200.flash_checkpoint.ddp import DdpCheckpointer, StorageType
        from oracle import oracle
        if rank == 0:
            AsyncCheckpointSaver.start_async_saving_ckpt()
        sd = _state(dev)
        want = [t.clone() for t in _tensors(sd)]

        def wreck():
            with torch.no_grad():
                for t in _tensors(sd):
                    t.fill_(3) if t.dtype != torch.bfloat16 else t.zero_()

        def check():
            for a, b in zip(_tensors(sd), want):
                assert torch.equal(a, b)

        # ---------------- striped save: one segment, two writers
        ckpt = DdpCheckpointer(tmpdir, stripe_replicas=True)
        handler = ckpt._engine._shm_handler
        for step in (1, 2):  # second save: cached layout, pinned stripes
            assert ckpt.save_checkpoint(step, sd, storage_type=StorageType.MEMORY)
        dist.barrier()
        _, total, expect = oracle.pack({"model_states": sd})
        got = np.frombuffer(handler.shared_memory.buf, dtype=np.uint8)
        assert handler.shared_memory.size == total
        assert got.tobytes() == expect.tobytes(), "striped image differs from the oracle's"
        del got
        entry = handler.metadata.get()[FCP_CHECKSUM_KEY]
        assert entry["sums"] == oracle.chunk_crc32c(expect, entry["chunk_bytes"]).tolist()
        step_s, lo, hi = handler.staged_image()
        assert step_s == 2 and 0 < hi - lo < total  # this rank holds its stripe only

        # ---------------- rank 1 lost its tensors; peers still hold their stripes
        if rank == 1:
            wreck()
        ckpt.load_checkpoint(into=sd)
        check()
        st = handler.last_stats
        assert st["restore_path"] == "slices_nvlink"
        assert st["bytes_over_host_link"] == 0 and st["peer_bytes_over_nvlink"] > 0

        # ---------------- restart: nothing in HBM, shm holds the image
        handler._drop_engine()
        wreck()
        ckpt.load_checkpoint(into=sd)
        check()
        st = handler.last_stats
        assert st["restore_path"] == "slices_nvlink"
        assert 0 < st["bytes_over_host_link"] < total  # my half only
        assert st["bytes_over_host_link"] + st["peer_bytes_over_nvlink"] == total

        # ---------------- a corrupt stripe fails the CRC on every rank
        if rank == 0:
            s0, lo0, _ = handler.staged_image()
            handler._engine.staging[lo0 + 11] ^= 1
        torch.cuda.synchronize()
        dist.barrier()
        with pytest.raises(RuntimeError, match="CRC-32C"):
            ckpt.load_checkpoint(into=sd)
        handler._drop_engine()  # forget the poisoned staging
        ckpt.load_checkpoint(into=sd)
        check()
        ckpt.close()
        dist.barrier()

        # ---------------- single-writer save (reference policy), then rank 1
        # recovers everything from rank 0's HBM
        ckpt2 = DdpCheckpointer(os.path.join(tmpdir, "b"))
        saved = ckpt2.save_checkpoint(5, sd, storage_type=StorageType.MEMORY)
        assert saved == (rank == 0)
        dist.barrier()
        h2 = ckpt2._engine._shm_handler
        if rank == 0:
            assert h2.staged_image() == (5, 0, total)
        else:
            wreck()
        ckpt2.load_checkpoint(into=sd)
        check()
        if rank == 1:
            assert h2.last_stats["bytes_over_host_link"] == 0
            assert h2.last_stats["peer_bytes_over_nvlink"] == total
        # fresh tensors on the device instead of `into`
        back = ckpt2.load_checkpoint(device=dev)
        for a, b in zip(_tensors(back), want):
            assert a.is_cuda and torch.equal(a, b)
        assert back["step"] == 9
        ckpt2.close()
        dist.barrier()
        if rank == 0:
            saver = AsyncCheckpointSaver.get_ckpt_saver()
            if saver:
                saver.close()
        dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        errq.put("rank %d:\n%s" % (rank, traceback.format_exc()))
        raise


@pytest.mark.timeout(300)
def test_two_rank_replicated_striped_save_and_nvlink_restore():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    ctx = mp.get_context("spawn")
    errq = ctx.SimpleQueue()
    port, run_id = _free_port(), "fcp_rep_%d" % os.getpid()
    with tempfile.TemporaryDirectory() as tmpdir:
        procs = [ctx.Process(target=_worker, args=(r, 2, port, run_id, tmpdir, errq))
                 for r in range(2)]
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
