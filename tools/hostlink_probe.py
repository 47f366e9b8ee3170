"""What can the host absorb?  Every rank copies a 4 GiB device buffer to / from
host memory at the same time, with the copy engine only (no kernels of this
repo), for three kinds of host memory: (a) cudaHostAlloc'ed, (b) a /dev/shm
mapping registered in one piece, (c) the same registered in 64 MiB pieces as
the engine does.  The aggregate over ranks is the host-side ceiling that
bench.py's save / restore rates at N GPUs have to be read against.
usage: torchrun --nproc-per-node N tools/hostlink_probe.py"""
import ctypes
import json
import mmap
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlrover_wuqiong_b200 import _native  # noqa: E402

GIB = 1 << 30
N = int(float(os.getenv("PROBE_GIB", "4")) * GIB)
PIECE = 64 << 20


def timed(fn, stream, reps=4):
    out = []
    for _ in range(reps):
        torch.cuda.synchronize()
        dist.barrier()
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record(stream)
        fn()
        b.record(stream)
        b.synchronize()
        t = torch.tensor([a.elapsed_time(b)], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out.append(t.item() / 1e3)
    return min(out[1:])


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.init()
    rt = ctypes.CDLL("libcudart.so.12")  # the runtime torch has already loaded
    rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t,
                                   ctypes.c_int, ctypes.c_void_p]
    stream = torch.cuda.current_stream()
    dev = torch.empty(N, dtype=torch.uint8, device="cuda")
    dev.random_(0, 256)
    res = {}

    def copy(dst, src, n, kind):
        err = rt.cudaMemcpyAsync(dst, src, n, kind, stream.cuda_stream)
        assert int(err) == 0, err

    D2H, H2D = 2, 1
    # (a) cudaHostAlloc
    pinned = torch.empty(N, dtype=torch.uint8).pin_memory()
    res["hostalloc_d2h"] = timed(lambda: copy(pinned.data_ptr(), dev.data_ptr(), N, D2H), stream)
    res["hostalloc_h2d"] = timed(lambda: copy(dev.data_ptr(), pinned.data_ptr(), N, H2D), stream)
    del pinned
    # (b), (c) /dev/shm mapping
    name = "/dev/shm/fcp_hostlink_probe_%d" % rank
    fd = os.open(name, os.O_CREAT | os.O_RDWR | os.O_TRUNC, 0o600)
    os.ftruncate(fd, N)
    mm = mmap.mmap(fd, N, flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)
    addr = ctypes.addressof(ctypes.c_char.from_buffer(mm))
    for label, piece in (("shm_one_piece", N), ("shm_64MiB_pieces", PIECE)):
        offs = list(range(0, N, piece))
        for o in offs:
            _native.host_register(addr + o, min(piece, N - o))

        def d2h():
            for o in offs:
                copy(addr + o, dev.data_ptr() + o, min(piece, N - o), D2H)

        def h2d():
            for o in offs:
                copy(dev.data_ptr() + o, addr + o, min(piece, N - o), H2D)

        res[label + "_d2h"] = timed(d2h, stream)
        res[label + "_h2d"] = timed(h2d, stream)
        torch.cuda.synchronize()
        for o in offs:
            _native.host_unregister(addr + o)
    mm.close()
    os.close(fd)
    os.unlink(name)
    if rank == 0:
        line = {"n_gpus": world, "gib_per_rank": N / GIB,
                "GBps_per_rank": {k: round(N / v / 1e9, 2) for k, v in res.items()},
                "GBps_aggregate": {k: round(world * N / v / 1e9, 1) for k, v in res.items()}}
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
