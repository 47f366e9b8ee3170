"""Config 5 of BASELINE.json: state_dict shape sweep — N tensors x size each,
pack / unpack + checksum through the C ABI against the reference's per-tensor
torch copy_ loop (oracle port) on the same CUDA tensors.  Prints a markdown
table; every cell is also parity-checked (CRC of the packed image vs the
oracle's CRC over the loop's output)."""
import mmap
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlrover_wuqiong_b200 import _native  # noqa: E402
from oracle import oracle  # noqa: E402

CAP = int(float(os.environ.get("SWEEP_CAP_GB", "8")) * 1e9)
NS = [10, 100, 1000, 10000]
SIZES = [4 << 10, 64 << 10, 1 << 20, 16 << 20, 256 << 20, 1 << 30]


def build(n, size, adamw_like):
    """mixed bf16 / fp32 tensors of `size` bytes; adamw_like puts a 4-byte
    scalar before each to force misalignment"""
    sd = {}
    for i in range(n):
        if adamw_like:
            sd["s%d" % i] = torch.tensor(float(i), device="cuda")
        dt = torch.bfloat16 if i % 2 else torch.float32
        es = 2 if i % 2 else 4
        sd["t%d" % i] = torch.empty(size // es, dtype=dt, device="cuda").normal_()
    return sd


def timed_events(fn, iters=3):
    fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(iters):
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def main():
    stream = lambda: torch.cuda.current_stream().cuda_stream  # noqa: E731
    print("| tensors | size each | layout | B (MB) | pack+CRC ms | pack GB/s (2B) | "
          "unpack+CRC ms | ref loop D2H ms | ref loop GB/s | speed-up |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    for n in NS:
        for size in SIZES:
            if n * size > CAP:
                continue
            for adamw in (False, True):
                if adamw and size not in (4 << 10, 1 << 20, 256 << 20):
                    continue
                sd = build(n, size, adamw)
                flat, meta, total = oracle.flatten(sd)
                descs = _native.make_descs([t.data_ptr() for t, m in flat],
                                           [m.offset for t, m in flat],
                                           [m.numel * m.element_size for t, m in flat])
                plan = _native.Plan(descs, total, 0)
                staging = torch.empty(plan.staging_bytes, dtype=torch.uint8, device="cuda")
                t_pack = timed_events(lambda: plan.pack(staging.data_ptr(), 1, stream()))
                sums = plan.read_checksums(stream())
                t_unpack = timed_events(lambda: plan.scatter(staging.data_ptr(), 1, stream()))
                # reference loop: per-tensor copy_ into a pageable mapping
                mm = mmap.mmap(-1, total, flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)
                oracle.ref_save_into(sd, meta, mm)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                oracle.ref_save_into(sd, meta, mm)
                torch.cuda.synchronize()
                t_ref = (time.perf_counter() - t0) * 1e3
                if total <= (512 << 20):  # parity of the cell
                    ref_sums = oracle.chunk_crc32c(np.frombuffer(mm, dtype=np.uint8),
                                                   _native.DEFAULT_CHUNK_BYTES)
                    assert ref_sums.tolist() == sums.tolist(), (n, size, adamw)
                mm.close()
                print("| %d | %s | %s | %.1f | %.3f | %.0f | %.3f | %.2f | %.2f | %.0fx |" % (
                    n, ("%d KiB" % (size >> 10)) if size < (1 << 20) else "%d MiB" % (size >> 20),
                    "4B scalar before each" if adamw else "dense", total / 1e6, t_pack,
                    2 * total / t_pack / 1e6, t_unpack, t_ref, total / t_ref / 1e6,
                    t_ref / t_pack), flush=True)
                plan.close()
                del sd, flat, staging
                torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
