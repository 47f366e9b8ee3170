"""Config 5's small-tensor corner: the cells of tools/sweep.py where launch
and per-tensor costs dominate (N x 4 KiB, 1000 x 64 KiB), pack+CRC and
scatter+CRC through the C ABI, CUDA events (median of 20).  Used before/after
kernel changes and under ncu (--ncu runs each call once, no timing loop).

    python tools/small_probe.py [--ncu]
"""
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlrover_wuqiong_b200 import _native  # noqa: E402
from oracle import oracle  # noqa: E402
from tools.sweep import build  # noqa: E402

if os.environ.get("FCP_LIB"):  # a differently built library (tuning experiments)
    _native.LIB_PATH = os.environ["FCP_LIB"]

CELLS = [(10, 4 << 10, False), (100, 4 << 10, False), (1000, 4 << 10, False),
         (1000, 4 << 10, True), (10000, 4 << 10, False), (10000, 4 << 10, True),
         (1000, 64 << 10, False), (100, 1 << 20, True)]


def med_ms(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    out = []
    for _ in range(iters):
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        out.append(a.elapsed_time(b))
    return statistics.median(out)


def gpu_ms(fn, reps=30):
    """GPU-side time per call: the calls are queued behind a long-running
    kernel so the host runs ahead and launch overhead drops out (a small image
    is otherwise bound by the ~30 us the host needs to enqueue one call)."""
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    torch.cuda._sleep(int(2e7))  # ~10 ms of spinning
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    b.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ncu = "--ncu" in sys.argv
    st = lambda: torch.cuda.current_stream().cuda_stream  # noqa: E731
    print("| tensors | size | layout | B (MB) | pack+CRC ms | GB/s (2B) | scatter+CRC ms | GB/s (2B) "
          "| pack GPU-only ms | scatter GPU-only ms |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    cells = CELLS
    for arg in sys.argv[1:]:
        if arg.startswith("--cells="):  # e.g. --cells=4,6 (indices into CELLS)
            cells = [CELLS[int(i)] for i in arg.split("=", 1)[1].split(",")]
    for n, size, adamw in cells:
        sd = build(n, size, adamw)
        flat, meta, total = oracle.flatten(sd)
        descs = _native.make_descs([t.data_ptr() for t, m in flat], [m.offset for t, m in flat],
                                   [m.numel * m.element_size for t, m in flat])
        plan = _native.Plan(descs, total, 0)
        staging = torch.empty(plan.staging_bytes, dtype=torch.uint8, device="cuda")
        plan.pack(staging.data_ptr(), 1, st())
        sums = plan.read_checksums(st())
        _, _, expect = oracle.pack(sd)
        assert staging[:total].cpu().numpy().tobytes() == expect.tobytes()
        assert sums.tolist() == oracle.chunk_crc32c(expect, _native.DEFAULT_CHUNK_BYTES).tolist()
        if ncu:
            plan.pack(staging.data_ptr(), 1, st())
            plan.scatter(staging.data_ptr(), 1, st())
            torch.cuda.synchronize()
            continue
        tp = med_ms(lambda: plan.pack(staging.data_ptr(), 1, st()))
        tu = med_ms(lambda: plan.scatter(staging.data_ptr(), 1, st()))
        gp = gpu_ms(lambda: plan.pack(staging.data_ptr(), 1, st()))
        gu = gpu_ms(lambda: plan.scatter(staging.data_ptr(), 1, st()))
        fresh = [torch.zeros_like(t) for t, m in flat]
        plan.update_ptrs([t.data_ptr() for t in fresh])
        plan.scatter(staging.data_ptr(), 1, st())
        assert plan.read_checksums(st()).tolist() == sums.tolist()
        for (a, _), b in zip(flat, fresh):
            assert torch.equal(a, b)
        print("| %d | %d KiB | %s | %.1f | %.4f | %.0f | %.4f | %.0f | %.4f | %.4f |" % (
            n, size >> 10, "4B scalar before each" if adamw else "dense", total / 1e6, tp,
            2 * total / tp / 1e6, tu, 2 * total / tu / 1e6, gp, gu), flush=True)
        plan.close()


if __name__ == "__main__":
    main()
