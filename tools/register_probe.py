"""How fast can a fresh process pin an existing shm image?  cudaHostRegister in
1..8 threads over disjoint slices (ctypes releases the GIL)."""
import mmap
import sys
import os
import threading
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlrover_wuqiong_b200 import _native  # noqa: E402

GB = 1 << 30
total = int(float(os.environ.get("PROBE_GB", "8")) * GB)
torch.cuda.init()
_native.lib().fcp_init(0)
mm = mmap.mmap(-1, total, flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)
arr = np.frombuffer(mm, dtype=np.uint8)
base = arr.ctypes.data
for nthreads in (1, 2, 4, 8):
    for slice_mb in (64, 256):
        sl = slice_mb << 20
        offs = list(range(0, total, sl))
        lock = threading.Lock()

        def work():
            while True:
                with lock:
                    if not offs:
                        return
                    o = offs.pop(0)
                _native.host_register(base + o, min(sl, total - o))

        t0 = time.time()
        ths = [threading.Thread(target=work) for _ in range(nthreads)]
        [t.start() for t in ths]
        [t.join() for t in ths]
        dt = time.time() - t0
        for o in range(0, total, sl):
            _native.host_unregister(base + o)
        print("threads=%d slice=%dMiB register %.3fs  %.1f GB/s" % (nthreads, slice_mb, dt, total / dt / 1e9), flush=True)
