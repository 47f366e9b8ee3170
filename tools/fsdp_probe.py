"""Where does an FSDP sharded save_to_memory spend its time?  2 ranks, gloo,
CPU: state-dict assembly (framework) vs DCP planning + collectives vs the byte
path.  usage: python tools/fsdp_probe.py [n_layer] [width]"""
import cProfile
import io
import os
import pstats
import sys
import tempfile
import time

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def worker(rank, world, port, run_id, n_layer, width, tmpdir):
    os.environ.update({
        "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
        "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank), "LOCAL_WORLD_SIZE": str(world),
        "GROUP_RANK": "0", "GROUP_WORLD_SIZE": "1", "NODE_NUM": "1", "NODE_RANK": "0",
        "ROLE_NAME": "dlrover-trainer", "TORCHELASTIC_RUN_ID": run_id,
        "DLROVER_LOG_LEVEL": "WARNING"})
    cuda = os.getenv("FSDP_PROBE_CUDA") == "1"
    if cuda:
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world,
                                device_id=torch.device("cuda", rank))
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    from torch.distributed.fsdp import FullyShardedDataParallel as FSDP
    from torch.distributed.fsdp import StateDictType
    from test_dist_cpu import TinyGPT
    from dlrover_wuqiong_b200.ckpt_saver import AsyncCheckpointSaver
    from dlrover_wuqiong_b200.common.constants import CheckpointConstant
    from dlrover_wuqiong_b200.flash_checkpoint.fsdp import FsdpCheckpointer
    if rank == 0:
        AsyncCheckpointSaver.start_async_saving_ckpt()
    torch.manual_seed(3)
    if cuda:
        import torch.nn as nn
        blocks = []
        for _ in range(n_layer):
            blocks += [nn.LayerNorm(width), nn.Linear(width, 4 * width), nn.GELU(),
                       nn.Linear(4 * width, width)]
        model = FSDP(nn.Sequential(*blocks).cuda(), device_id=torch.device("cuda", rank))
        opt = torch.optim.AdamW(model.parameters(), lr=1e-4)
        model(torch.randn(2, width, device="cuda")).square().mean().backward()
    else:
        model = FSDP(TinyGPT(n_layer=n_layer, e=width, vocab=1001, block=16),
                     device_id=torch.device("cpu"))
        opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
        model(torch.randint(0, 1001, (2, 16))).sum().backward()
    opt.step()
    ckpt = FsdpCheckpointer(tmpdir, async_save=os.getenv("FSDP_PROBE_ASYNC") == "1")
    eng = ckpt._engine
    paths = {CheckpointConstant.MODEL_STATES_NAME: os.path.join(tmpdir, "x")}
    rows = []
    for step in range(1, 7):
        if cuda:
            torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        with FsdpCheckpointer._sharded_state(model):
            sd = {"model": model.state_dict(), "optim": FSDP.optim_state_dict(model, opt), "step": step}
        t1 = time.perf_counter()
        prof = cProfile.Profile() if (rank == 0 and step == 5) else None
        if prof:
            prof.enable()
        ok = eng.save_to_memory(step, sd, paths)
        if prof:
            prof.disable()
            s = io.StringIO()
            pstats.Stats(prof, stream=s).sort_stats("cumulative").print_stats(45)
            print(s.getvalue()[:9000])
        t2 = time.perf_counter()
        assert ok
        rows.append((t1 - t0, t2 - t1, eng.last_save_planned))
    if rank == 0:
        h = eng._shm_handler
        n_items = len(h.metadata.get()["dcp_metadata"].storage_data)
        print("items in the global plan: %d, this rank's segment: %.1f MB" % (
            n_items, h.shared_memory.size / 1e6))
        for i, (a, b, planned) in enumerate(rows, 1):
            print("save %d: state-dict assembly %.1f ms, save_to_memory %.1f ms (%s)" % (i, a * 1e3, b * 1e3, "planned" if planned else "cached plan"))
    # ---- restore from memory, profiled on rank 0
    ckpt.wait_latest_checkpoint()
    for rep in range(2):
        if cuda:
            torch.cuda.synchronize()
        dist.barrier()
        prof = cProfile.Profile() if (rank == 0 and rep == 1) else None
        t0 = time.perf_counter()
        if prof:
            prof.enable()
        extra = ckpt.load_checkpoint(model, opt)
        if cuda:
            torch.cuda.synchronize()
        if prof:
            prof.disable()
            s = io.StringIO()
            pstats.Stats(prof, stream=s).sort_stats("cumulative").print_stats(60)
            print(s.getvalue()[:14000])
        if rank == 0:
            print("load %d: load_checkpoint %.1f ms (step %s)" % (
                rep, (time.perf_counter() - t0) * 1e3, extra.get("step")))
    dist.barrier()
    ckpt.close()
    if rank == 0:
        AsyncCheckpointSaver.get_ckpt_saver().close()
    dist.destroy_process_group()


if __name__ == "__main__":
    n_layer = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    width = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(worker, args=(2, port, "fcp_fsdp_probe_%d" % os.getpid(), n_layer, width, d),
                 nprocs=2, join=True)
