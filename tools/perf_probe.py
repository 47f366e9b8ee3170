"""Quick device-side probe of the raw pieces (not the bench): pack / scatter
kernel rates with and without the fused CRC, aligned and 4-byte-shifted
layouts, and the host-link rates of the staged and direct modes."""
import json
import mmap
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlrover_wuqiong_b200 import _native  # noqa: E402

GB = 1e9


def build(total_gb, n_tensors, shift):
    dev = torch.device("cuda", 0)
    per = int(total_gb * GB / n_tensors) // 16 * 16
    tensors, offs, off = [], [], 0
    for i in range(n_tensors):
        t = torch.empty(per, dtype=torch.uint8, device=dev)
        t.random_(0, 256)
        tensors.append(t)
        off += shift  # a `shift`-byte hole before every tensor (like a step scalar)
        offs.append(off)
        off += per
    d = _native.make_descs([t.data_ptr() for t in tensors], offs, [per] * n_tensors)
    return tensors, d, off


def timed(fn, iters=5, warm=2):
    st = torch.cuda.current_stream()
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record(st)
        fn()
        b.record(st)
        b.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts), sorted(ts)[len(ts) // 2]


def main():
    total_gb = float(os.environ.get("PROBE_GB", "8"))
    out = {}
    s = lambda: torch.cuda.current_stream().cuda_stream  # noqa: E731
    for shift in ((0, 4) if os.environ.get('PROBE_QUICK') else (0, 4, 2)):
        tensors, d, total = build(total_gb, 64, shift)
        plan = _native.Plan(d, total, 0)
        staging = torch.empty(plan.staging_bytes, dtype=torch.uint8, device="cuda")
        for name, cs in (("crc", _native.SUM_CRC32C), ("nocrc", _native.SUM_NONE)):
            best, med = timed(lambda: plan.pack(staging.data_ptr(), cs, s()))
            out["pack_%s_shift%d" % (name, shift)] = {
                "ms": best, "ms_med": med, "GBps_2B": 2 * total / best / 1e6}
            best, med = timed(lambda: plan.scatter(staging.data_ptr(), cs, s()))
            out["scatter_%s_shift%d" % (name, shift)] = {
                "ms": best, "ms_med": med, "GBps_2B": 2 * total / best / 1e6}
        if shift == 0:
            xo = torch.empty(plan.num_chunks, dtype=torch.int64, device="cuda")
            best, med = timed(lambda: _native.chunk_xxh64(staging.data_ptr(), total,
                                                          xo.data_ptr(), stream=s()))
            out["xxh64_pass"] = {"ms": best, "ms_med": med, "GBps_1B": total / best / 1e6}
        if shift == 4 and not os.environ.get('PROBE_QUICK'):
            # host link
            mm = mmap.mmap(-1, total, flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)
            harr = np.frombuffer(mm, dtype=np.uint8)
            t0 = time.time()
            _native.host_register(harr.ctypes.data, total)
            out["host_register_s_per_GB"] = (time.time() - t0) / (total / GB)
            for mode, mname in ((_native.MODE_STAGED, "staged"), (_native.MODE_DIRECT, "direct")):
                def save():
                    plan.save_async(staging.data_ptr(), harr.ctypes.data, mode,
                                    _native.SUM_CRC32C, s())
                    plan.save_wait()
                ts = []
                for _ in range(4):
                    torch.cuda.synchronize()
                    t0 = time.time()
                    save()
                    ts.append(time.time() - t0)
                tm = plan.timings()
                out["save_%s" % mname] = {
                    "wall_s": min(ts[1:]), "GBps": total / min(ts[1:]) / GB, **tm}
                sums = plan.save_wait  # noqa
                ts = []
                for _ in range(4):
                    torch.cuda.synchronize()
                    t0 = time.time()
                    plan.restore(harr.ctypes.data, staging.data_ptr(), mode,
                                 _native.SUM_CRC32C, None, s())
                    ts.append(time.time() - t0)
                out["restore_%s" % mname] = {
                    "wall_s": min(ts[1:]), "GBps": total / min(ts[1:]) / GB}
            torch.cuda.synchronize()
            _native.host_unregister(harr.ctypes.data)
            del harr
            mm.close()
        plan.close()
        del tensors, staging
        torch.cuda.empty_cache()
    # plain torch D2D copy for reference (same 2B accounting)
    a = torch.empty(int(total_gb * GB), dtype=torch.uint8, device="cuda")
    b = torch.empty_like(a)
    best, _ = timed(lambda: b.copy_(a))
    out["torch_copy"] = {"ms": best, "GBps_2B": 2 * a.numel() / best / 1e6}
    if os.environ.get('PROBE_QUICK'):
        for k, v in out.items():
            print(k, {a: round(b, 2) for a, b in v.items()})
    else:
        print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
