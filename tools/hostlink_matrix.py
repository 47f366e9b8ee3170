"""Can the host-side ceiling of concurrent saves be moved from inside the guest?
One torchrun launch, every rank one GPU, copy engine only unless said
otherwise; prints one JSON line per experiment on rank 0.

  pages    device->host into (a) cudaHostAlloc memory, (b) a registered /dev/shm
           mapping (4 KiB pages), (c) the same after madvise(MADV_HUGEPAGE),
           (d) anonymous memory with MADV_HUGEPAGE (THP), registered,
           (e) MAP_HUGETLB anonymous memory, registered — all ranks at once
  writers  k = 1, 2, 4, ... ranks copy at a time, in world/k turns, same total
           bytes: does staggering beat everybody at once?
  sizes    one cudaMemcpyAsync per 4 MiB ... 1 GiB piece, all ranks at once
  sm       (--sm) the pack kernel's staged_kernel leg: SM stores into the
           mapped segment with FCP_COPY_CTAS CTAs (set per launch)

usage: torchrun --nproc-per-node N tools/hostlink_matrix.py [--gib 2] [--sm]
"""
import argparse
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
MADV_HUGEPAGE, MADV_POPULATE_WRITE = 14, 23
MAP_HUGETLB = 0x40000


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gib", type=float, default=2.0)
    ap.add_argument("--sm", action="store_true")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    N = int(args.gib * GIB)
    rt = ctypes.CDLL("libcudart.so.12")
    rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t,
                                   ctypes.c_int, ctypes.c_void_p]
    libc = ctypes.CDLL(None, use_errno=True)
    libc.madvise.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
    stream = torch.cuda.current_stream()
    dev = torch.empty(N, dtype=torch.uint8, device="cuda")
    dev.random_(0, 256)
    D2H, H2D = 2, 1

    def say(**kw):
        if rank == 0:
            print(json.dumps(kw), flush=True)

    def copy(dst, src, n, kind):
        err = rt.cudaMemcpyAsync(dst, src, n, kind, stream.cuda_stream)
        assert int(err) == 0, err

    def timed(fn, active=True, reps=3):
        """max over ranks of the device time of fn() (ranks with active=False
        only take part in the barriers); best of reps-1."""
        out = []
        for _ in range(reps):
            torch.cuda.synchronize()
            dist.barrier()
            a, b = torch.cuda.Event(True), torch.cuda.Event(True)
            a.record(stream)
            if active:
                fn()
            b.record(stream)
            b.synchronize()
            t = torch.tensor([a.elapsed_time(b) if active else 0.0], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            out.append(t.item() / 1e3)
        return min(out[1:])

    def both_ways(addr, piece=None):
        piece = piece or N
        offs = list(range(0, N, piece))

        def d2h():
            for o in offs:
                copy(addr + o, dev.data_ptr() + o, min(piece, N - o), D2H)

        def h2d():
            for o in offs:
                copy(dev.data_ptr() + o, addr + o, min(piece, N - o), H2D)

        return timed(d2h), timed(h2d)

    def report(what, label, td, th, **extra):
        say(experiment=what, case=label, n_gpus=world, gib_per_rank=args.gib,
            d2h_GBps_per_rank=round(N / td / 1e9, 2), h2d_GBps_per_rank=round(N / th / 1e9, 2),
            d2h_GBps_aggregate=round(world * N / td / 1e9, 1),
            h2d_GBps_aggregate=round(world * N / th / 1e9, 1), **extra)

    def thp_state():
        out = {}
        for k in ("enabled", "shmem_enabled", "defrag"):
            try:
                out[k] = open("/sys/kernel/mm/transparent_hugepage/" + k).read().strip()
            except OSError:
                out[k] = None
        try:
            out["nr_hugepages"] = int(open("/proc/sys/vm/nr_hugepages").read())
        except OSError:
            out["nr_hugepages"] = None
        return out

    def huge_kb(addr_lo, addr_hi):
        """AnonHugePages + ShmemPmdMapped of the mapping at addr_lo (smaps)."""
        try:
            kb, inside = 0, False
            for line in open("/proc/self/smaps"):
                if "-" in line.split(" ")[0] and line[0] in "0123456789abcdef":
                    lo, hi = (int(x, 16) for x in line.split(" ")[0].split("-"))
                    inside = lo <= addr_lo < hi
                elif inside and line.split(":")[0] in ("AnonHugePages", "ShmemPmdMapped"):
                    kb += int(line.split()[1])
            return kb
        except Exception:  # noqa: BLE001
            return None

    say(experiment="host", thp=thp_state(), cpus=os.cpu_count())

    if not args.sm:
        # ---------------------------------------------------------------- pages
        pinned = torch.empty(N, dtype=torch.uint8).pin_memory()
        report("pages", "cudaHostAlloc", *both_ways(pinned.data_ptr()))
        # ------------------------------------------------------------ writers
        k = 1
        while k <= world:
            turns = world // k
            total = 0.0
            for turn in range(turns):
                mine = rank // k == turn
                total += timed(lambda: copy(pinned.data_ptr(), dev.data_ptr(), N, D2H), mine)
            say(experiment="writers", concurrent=k, turns=turns, n_gpus=world,
                d2h_GBps_aggregate=round(world * N / total / 1e9, 1),
                d2h_GBps_per_active_rank=round(N / (total / turns) / 1e9, 2))
            total = 0.0
            for turn in range(turns):
                mine = rank // k == turn
                total += timed(lambda: copy(dev.data_ptr(), pinned.data_ptr(), N, H2D), mine)
            say(experiment="readers", concurrent=k, turns=turns, n_gpus=world,
                h2d_GBps_aggregate=round(world * N / total / 1e9, 1),
                h2d_GBps_per_active_rank=round(N / (total / turns) / 1e9, 2))
            k *= 2
        # -------------------------------------------------------------- sizes
        for piece in (4 << 20, 16 << 20, 64 << 20, 256 << 20, 1 << 30):
            if piece <= N:
                report("sizes", "%d MiB per copy" % (piece >> 20),
                       *both_ways(pinned.data_ptr(), piece))
        del pinned

        def registered(label, mm, populate=True, advise=False):
            addr = ctypes.addressof(ctypes.c_char.from_buffer(mm))
            adv = libc.madvise(addr, N, MADV_HUGEPAGE) if advise else None
            if populate:
                if libc.madvise(addr, N, MADV_POPULATE_WRITE) != 0:
                    ctypes.memset(addr, 1, N)
            try:
                _native.host_register(addr, N)
            except Exception as e:  # noqa: BLE001
                say(experiment="pages", case=label, unavailable=str(e)[:120])
                return
            td, th = both_ways(addr)
            report("pages", label, td, th, madvise_rc=adv, huge_mapped_kb=huge_kb(addr, addr + N))
            torch.cuda.synchronize()
            _native.host_unregister(addr)

        name = "/dev/shm/fcp_hostlink_matrix_%d" % rank
        for label, advise in (("shm, 4 KiB pages", False), ("shm + MADV_HUGEPAGE", True)):
            fd = os.open(name, os.O_CREAT | os.O_RDWR | os.O_TRUNC, 0o600)
            os.ftruncate(fd, N)
            mm = mmap.mmap(fd, N, flags=mmap.MAP_SHARED)
            registered(label, mm, advise=advise)
            mm.close()
            os.close(fd)
            os.unlink(name)
        mm = mmap.mmap(-1, N, flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
        registered("anonymous + MADV_HUGEPAGE (THP)", mm, advise=True)
        mm.close()
        try:
            mm = mmap.mmap(-1, N, flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS | MAP_HUGETLB)
            registered("MAP_HUGETLB", mm)
            mm.close()
        except OSError as e:
            say(experiment="pages", case="MAP_HUGETLB", unavailable=str(e))
    else:
        # ----------------------------------------------- SM stores into the host
        ctas = os.getenv("FCP_COPY_CTAS", "32")
        name = "/dev/shm/fcp_hostlink_matrix_%d" % rank
        fd = os.open(name, os.O_CREAT | os.O_RDWR | os.O_TRUNC, 0o600)
        os.ftruncate(fd, N)
        mm = mmap.mmap(fd, N, flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)
        addr = ctypes.addressof(ctypes.c_char.from_buffer(mm))
        piece = 64 << 20
        for o in range(0, N, piece):
            _native.host_register(addr + o, min(piece, N - o))
        plan = _native.Plan(_native.make_descs([dev.data_ptr()], [0], [N]), N, local)
        plan.set_copy_piece_bytes(piece)
        staging = torch.empty(plan.staging_bytes, dtype=torch.uint8, device="cuda")
        res = {}
        for label, mode in (("copy_engine", _native.MODE_STAGED),
                            ("sm_stores", _native.MODE_STAGED_KERNEL)):
            def save():
                plan.save_async(staging.data_ptr(), addr, mode, _native.SUM_CRC32C,
                                stream.cuda_stream)
                plan.save_wait()
            res[label] = timed(save)
        say(experiment="sm", copy_ctas=int(ctas), n_gpus=world, gib_per_rank=args.gib,
            copy_engine_GBps_per_rank=round(N / res["copy_engine"] / 1e9, 2),
            sm_stores_GBps_per_rank=round(N / res["sm_stores"] / 1e9, 2),
            copy_engine_GBps_aggregate=round(world * N / res["copy_engine"] / 1e9, 1),
            sm_stores_GBps_aggregate=round(world * N / res["sm_stores"] / 1e9, 1))
        plan.close()
        torch.cuda.synchronize()
        for o in range(0, N, piece):
            _native.host_unregister(addr + o)
        mm.close()
        os.close(fd)
        os.unlink(name)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
