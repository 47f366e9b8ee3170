"""Does a D2H copy into a range pinned piece by piece run at full speed?"""
import mmap, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlrover_wuqiong_b200 import _native
GB = 1e9
total_gb = 8
per = int(total_gb * GB / 32) // 32 * 32
tensors = [torch.empty(per, dtype=torch.uint8, device="cuda").random_(0, 256) for _ in range(32)]
total = per * 32
plan = _native.Plan(_native.make_descs([t.data_ptr() for t in tensors], [i * per for i in range(32)], [per] * 32), total, 0)
staging = torch.empty(plan.staging_bytes, dtype=torch.uint8, device="cuda")
mm = mmap.mmap(-1, total, flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)
arr = np.frombuffer(mm, dtype=np.uint8)
s = torch.cuda.current_stream().cuda_stream
for piece_mb in (0, 128, 64):
    if piece_mb == 0:
        _native.host_register(arr.ctypes.data, total); pieces = [0]
    else:
        p = piece_mb << 20
        pieces = list(range(0, total, p))
        for o in pieces:
            _native.host_register(arr.ctypes.data + o, min(p, total - o))
    ts = []
    for _ in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        plan.save_async(staging.data_ptr(), arr.ctypes.data, 0, 1, s); sums = plan.save_wait()
        ts.append(time.perf_counter() - t0)
    ok = bytes(arr[:64]) == bytes(tensors[0][:64].cpu().numpy())
    t0 = time.perf_counter(); plan.restore(arr.ctypes.data, staging.data_ptr(), 0, 1, sums, s); tr = time.perf_counter() - t0
    print("pieces of %4d MiB: save %.1f GB/s restore %.1f GB/s ok=%s" % (piece_mb, total / min(ts) / GB, total / tr / GB, ok), flush=True)
    for o in pieces:
        _native.host_unregister(arr.ctypes.data + o)
