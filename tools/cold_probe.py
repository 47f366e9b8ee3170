"""Bounced (unregistered) restore rate vs image size and worker count."""
import mmap, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlrover_wuqiong_b200 import _native
GB = 1e9
for total_gb in [float(x) for x in os.getenv('COLD_GB', '4,12,24').split(',')]:
    per = int(total_gb * GB / 32) // 32 * 32
    tensors = [torch.empty(per, dtype=torch.uint8, device="cuda") for _ in range(32)]
    offs = [i * per for i in range(32)]
    total = per * 32
    plan = _native.Plan(_native.make_descs([t.data_ptr() for t in tensors], offs, [per] * 32), total, 0)
    staging = torch.empty(plan.staging_bytes, dtype=torch.uint8, device="cuda")
    mm = mmap.mmap(-1, total, flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)
    arr = np.frombuffer(mm, dtype=np.uint8)
    arr[::4096] = 7
    s = torch.cuda.current_stream().cuda_stream
    for workers in [int(x) for x in os.getenv('COLD_WORKERS', '4,8,12,16').split(',')]:
        ts = []
        for _ in range(3):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            plan.restore_bounced(arr.ctypes.data, staging.data_ptr(), 1, None, workers, s)
            ts.append(time.perf_counter() - t0)
        print("slot %s MiB size %2d GB workers %2d: first %.3fs best %.3fs  %.1f GB/s" % (os.getenv("FCP_BOUNCE_SLOT_MB", "8"), total_gb, workers, ts[0], min(ts), total / min(ts) / GB), flush=True)
    plan.close(); del tensors, staging, arr; mm.close(); torch.cuda.empty_cache()
