"""How fast can a fresh /dev/shm segment be made resident?  MAP_POPULATE (what
SharedMemory(create=True) does) vs madvise(MADV_POPULATE_WRITE) from several
threads vs fallocate from several threads followed by a mapping.  CPU only.
usage: python tools/populate_probe.py [GiB]"""
import ctypes
import mmap
import os
import sys
import threading
import time

libc = ctypes.CDLL("libc.so.6", use_errno=True)
libc.madvise.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
MADV_POPULATE_WRITE = 23
NAME = "/dev/shm/fcp_populate_probe"


def _threads(n, fn):
    ts = [threading.Thread(target=fn, args=(i,)) for i in range(n)]
    [t.start() for t in ts]
    [t.join() for t in ts]


def run(nbytes, threads, how):
    fd = os.open(NAME, os.O_CREAT | os.O_RDWR | os.O_TRUNC, 0o600)
    os.ftruncate(fd, nbytes)
    per = (nbytes // threads + 4095) // 4096 * 4096
    errs = []
    t0 = time.perf_counter()
    if how == "map_populate":
        mm = mmap.mmap(fd, nbytes, flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)
    elif how == "madvise":
        mm = mmap.mmap(fd, nbytes, flags=mmap.MAP_SHARED)
        addr = ctypes.addressof(ctypes.c_char.from_buffer(mm))

        def work(i):
            lo, hi = i * per, min(nbytes, (i + 1) * per)
            if hi > lo and libc.madvise(addr + lo, hi - lo, MADV_POPULATE_WRITE) != 0:
                errs.append(ctypes.get_errno())

        _threads(threads, work)
    else:  # fallocate in threads, then map the now-resident pages
        def work(i):
            lo, hi = i * per, min(nbytes, (i + 1) * per)
            if hi > lo:
                os.posix_fallocate(fd, lo, hi - lo)

        _threads(threads, work)
        t1 = time.perf_counter()
        mm = mmap.mmap(fd, nbytes, flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)
        errs.append("map %.3f s" % (time.perf_counter() - t1))
    dt = time.perf_counter() - t0
    mm.close()
    os.close(fd)
    os.unlink(NAME)
    return dt, errs


def main():
    n = int(float(sys.argv[1] if len(sys.argv) > 1 else 4) * (1 << 30))
    print("cpus", os.cpu_count())
    for how, th in (("map_populate", 1), ("madvise", 1), ("madvise", 4), ("madvise", 8),
                    ("madvise", 16), ("fallocate", 1), ("fallocate", 4), ("fallocate", 8)):
        dt, notes = run(n, th, how)
        print("%-13s threads=%2d  %.3f s  %.2f GB/s  %s" % (how, th, dt, n / 1e9 / dt, notes or ""))


if __name__ == "__main__":
    main()
