"""ctypes binding of libfcp.so (C ABI in include/fcp.h).

There is no CPU fallback: if the library is missing, ``lib()`` raises.  The
library is built in-tree by ``dlrover_wuqiong_b200.csrc.build`` so that the
``.so`` travels with the source tree.
"""
import ctypes
import errno
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libfcp.so")

REGION_BYTES = 16384
NOALLOC_MAX_BYTES = 65536 * REGION_BYTES
DEFAULT_CHUNK_BYTES = 2 << 20
SUM_NONE, SUM_CRC32C = 0, 1
MODE_STAGED, MODE_DIRECT, MODE_STAGED_KERNEL = 0, 1, 2

# struct fcp_desc { const void* ptr; uint64_t dst_off; uint64_t nbytes; }
DESC_DTYPE = np.dtype([("ptr", "<u8"), ("dst_off", "<u8"), ("nbytes", "<u8")])

# every symbol include/fcp.h declares: (name, restype, argtypes)
_vp, _u64, _u32, _i = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_int
SYMBOLS = [
    ("fcp_version", _i, []),
    ("fcp_last_error", ctypes.c_char_p, []),
    ("fcp_init", _i, [_i]),
    ("fcp_plan_create", _i, [_vp, _u64, _u64, _u32, _i, ctypes.POINTER(_vp)]),
    ("fcp_plan_update_ptrs", _i, [_vp, _vp, _u64]),
    ("fcp_plan_destroy", None, [_vp]),
    ("fcp_plan_total_bytes", _u64, [_vp]),
    ("fcp_plan_num_chunks", _u64, [_vp]),
    ("fcp_plan_num_regions", _u64, [_vp]),
    ("fcp_host_register", _i, [_vp, _u64, ctypes.POINTER(_vp)]),
    ("fcp_host_unregister", _i, [_vp]),
    ("fcp_host_chunk_crc32c", _i, [_vp, _u64, _u32, _vp, _i]),
    ("fcp_enable_peer", _i, [_i, _i]),
    ("fcp_ipc_export", _i, [_vp, _vp, _vp]),
    ("fcp_ipc_open", _i, [_vp, ctypes.POINTER(_vp)]),
    ("fcp_ipc_close", _i, [_vp]),
    ("fcp_pack", _i, [_vp, _vp, _i, _vp]),
    ("fcp_scatter", _i, [_vp, _vp, _i, _vp]),
    ("fcp_read_checksums", _i, [_vp, _vp, _vp]),
    ("fcp_launch", _i, [_vp, _u64, _u64, _u32, _u64, _u64, _vp, _vp, _vp, _i, _i, _vp]),
    ("fcp_finish_checksums", _i, [_vp, _u64, _u64, _u32, _vp]),
    ("fcp_chunk_xxh64", _i, [_vp, _u64, _u32, _vp, _vp]),
    ("fcp_launch_noalloc", _i, [_vp, _u64, _u64, _u32, _vp, _vp, _i, _i, _vp]),
    ("fcp_save_async", _i, [_vp, _vp, _vp, _i, _i, _vp]),
    ("fcp_plan_set_defer_copies", _i, [_vp, _i]),
    ("fcp_save_release_copies", _i, [_vp]),
    ("fcp_save_packed_wait", _i, [_vp]),
    ("fcp_save_wait", _i, [_vp, _vp]),
    ("fcp_restore", _i, [_vp, _vp, _vp, _i, _i, _vp, _vp]),
    ("fcp_restore_bounced", _i, [_vp, _vp, _vp, _i, _vp, _i, _vp]),
    ("fcp_restore_from_device", _i, [_vp, _vp, _i, _vp, _vp]),
    ("fcp_restore_from_slices", _i, [_vp, _vp, _u64, _u64, _i, _vp, _i, _u64, _vp]),
    ("fcp_stage_range", _i, [_vp, _vp, _vp, _u64, _u64, _i, _i, _vp]),
    ("fcp_plan_set_window", _i, [_vp, _u64, _u64]),
    ("fcp_plan_timings", _i, [_vp, _vp]),
    ("fcp_plan_last_launches", _u64, [_vp]),
    ("fcp_plan_set_copy_piece_bytes", _i, [_vp, _u64]),
    ("fcp_plan_set_read_prefetch", _i, [_vp, _i]),
    ("fcp_plan_set_slice_bytes", _i, [_vp, _u64]),
]

_lib = None


class FcpError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libfcp: %s (%s)" % (msg, errno.errorcode.get(-code, code)))
        self.code = code


class ChecksumMismatch(FcpError):
    pass


def lib():
    """Load libfcp.so, loudly."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "%s is missing: build it with `python -m "
                "dlrover_wuqiong_b200.csrc.build` (there is no CPU fallback "
                "for the flash-checkpoint GPU path)" % LIB_PATH)
        l = ctypes.CDLL(LIB_PATH)
        for name, res, args in SYMBOLS:
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc):
    if rc == 0:
        return
    msg = lib().fcp_last_error().decode("utf-8", "replace")
    if rc == -errno.EBADMSG:
        raise ChecksumMismatch(rc, msg)
    raise FcpError(rc, msg)


def make_descs(ptrs, offsets, nbytes):
    d = np.zeros(len(ptrs), dtype=DESC_DTYPE)
    d["ptr"] = ptrs
    d["dst_off"] = offsets
    d["nbytes"] = nbytes
    return d


class Plan:
    """Owner of one fcp_plan."""

    def __init__(self, descs, total_bytes, device, chunk_bytes=DEFAULT_CHUNK_BYTES):
        descs = np.ascontiguousarray(descs, dtype=DESC_DTYPE)
        self._h = ctypes.c_void_p()
        check(lib().fcp_plan_create(
            descs.ctypes.data if len(descs) else None, len(descs), total_bytes,
            chunk_bytes, device, ctypes.byref(self._h)))
        self.total_bytes = int(total_bytes)
        self.chunk_bytes = int(chunk_bytes)
        self.device = device
        self.n_desc = len(descs)
        self.num_chunks = int(lib().fcp_plan_num_chunks(self._h))
        self.num_regions = int(lib().fcp_plan_num_regions(self._h))

    @property
    def staging_bytes(self):
        return max(self.num_regions, 1) * REGION_BYTES

    def update_ptrs(self, ptrs):
        a = np.ascontiguousarray(ptrs, dtype=np.uint64)
        check(lib().fcp_plan_update_ptrs(self._h, a.ctypes.data if len(a) else None, len(a)))

    def set_slice_bytes(self, n):
        check(lib().fcp_plan_set_slice_bytes(self._h, n))

    def pack(self, packed_ptr, checksum=SUM_CRC32C, stream=0):
        check(lib().fcp_pack(self._h, packed_ptr, checksum, stream))

    def scatter(self, packed_ptr, checksum=SUM_CRC32C, stream=0):
        check(lib().fcp_scatter(self._h, packed_ptr, checksum, stream))

    def read_checksums(self, stream=0):
        out = np.zeros(self.num_chunks, dtype=np.uint32)
        check(lib().fcp_read_checksums(self._h, out.ctypes.data, stream))
        return out

    def save_async(self, staging_ptr, host_ptr, mode=MODE_STAGED,
                   checksum=SUM_CRC32C, stream=0):
        check(lib().fcp_save_async(self._h, staging_ptr, host_ptr, mode, checksum, stream))

    def set_defer_copies(self, defer):
        check(lib().fcp_plan_set_defer_copies(self._h, 1 if defer else 0))

    def release_copies(self):
        check(lib().fcp_save_release_copies(self._h))

    def save_packed_wait(self):
        check(lib().fcp_save_packed_wait(self._h))

    def save_wait(self, want_sums=True):
        out = np.zeros(self.num_chunks, dtype=np.uint32) if want_sums else None
        check(lib().fcp_save_wait(self._h, out.ctypes.data if want_sums else None))
        return out

    def restore(self, host_ptr, staging_ptr, mode=MODE_STAGED, checksum=SUM_CRC32C,
                expected=None, stream=0):
        exp = None
        if expected is not None:
            exp = np.ascontiguousarray(expected, dtype=np.uint32)
            if len(exp) != self.num_chunks:
                raise ValueError("expected %d checksums, got %d" % (self.num_chunks, len(exp)))
        check(lib().fcp_restore(self._h, host_ptr, staging_ptr, mode, checksum,
                                exp.ctypes.data if exp is not None else None, stream))

    def restore_bounced(self, host_ptr, staging_ptr, checksum=SUM_CRC32C, expected=None,
                        n_workers=0, stream=0):
        exp = None
        if expected is not None:
            exp = np.ascontiguousarray(expected, dtype=np.uint32)
            if len(exp) != self.num_chunks:
                raise ValueError("expected %d checksums, got %d" % (self.num_chunks, len(exp)))
        check(lib().fcp_restore_bounced(self._h, host_ptr, staging_ptr, checksum,
                                        exp.ctypes.data if exp is not None else None,
                                        n_workers, stream))

    def restore_from_device(self, packed_ptr, checksum=SUM_CRC32C, expected=None, stream=0):
        exp = None
        if expected is not None:
            exp = np.ascontiguousarray(expected, dtype=np.uint32)
        check(lib().fcp_restore_from_device(
            self._h, packed_ptr, checksum,
            exp.ctypes.data if exp is not None else None, stream))

    def restore_from_slices(self, bases, slice_bytes, checksum=SUM_CRC32C, expected=None,
                            read_prefetch=False, first_slice=0, stream=0):
        """bases[s]: address image byte 0 would have in the buffer that holds
        slice s (see fcp.h)."""
        a = np.ascontiguousarray(bases, dtype=np.uint64)
        exp = None
        if expected is not None:
            exp = np.ascontiguousarray(expected, dtype=np.uint32)
            if len(exp) != self.num_chunks:
                raise ValueError("expected %d checksums, got %d" % (self.num_chunks, len(exp)))
        check(lib().fcp_restore_from_slices(
            self._h, a.ctypes.data, len(a), slice_bytes, checksum,
            exp.ctypes.data if exp is not None else None, 1 if read_prefetch else 0,
            first_slice % max(len(a), 1), stream))

    def stage_range(self, host_ptr, staging_ptr, byte_begin, byte_end, registered=False,
                    n_workers=0, stream=0):
        check(lib().fcp_stage_range(self._h, host_ptr, staging_ptr, byte_begin, byte_end,
                                    1 if registered else 0, n_workers, stream))

    def set_window(self, chunk_begin=0, chunk_end=0xFFFFFFFFFFFFFFFF):
        check(lib().fcp_plan_set_window(self._h, chunk_begin, chunk_end))

    def timings(self):
        out = (ctypes.c_float * 3)()
        check(lib().fcp_plan_timings(self._h, out))
        return {"kernel_ms": out[0], "copy_ms": out[1], "total_ms": out[2]}

    def set_copy_piece_bytes(self, n):
        check(lib().fcp_plan_set_copy_piece_bytes(self._h, n))

    def set_read_prefetch(self, enable):
        check(lib().fcp_plan_set_read_prefetch(self._h, 1 if enable else 0))

    def last_launches(self):
        return int(lib().fcp_plan_last_launches(self._h))

    def close(self):
        if self._h:
            lib().fcp_plan_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def host_register(addr, nbytes):
    alias = ctypes.c_void_p()
    check(lib().fcp_host_register(addr, nbytes, ctypes.byref(alias)))
    return alias.value


def ipc_export(dev_ptr):
    """-> (handle bytes, offset of dev_ptr inside its allocation)"""
    handle = (ctypes.c_uint8 * 64)()
    offset = ctypes.c_uint64()
    check(lib().fcp_ipc_export(dev_ptr, handle, ctypes.byref(offset)))
    return bytes(handle), int(offset.value)


def ipc_open(handle):
    base = ctypes.c_void_p()
    buf = (ctypes.c_uint8 * 64).from_buffer_copy(handle)
    check(lib().fcp_ipc_open(buf, ctypes.byref(base)))
    return base.value


def ipc_close(base):
    check(lib().fcp_ipc_close(base))


class RawDeviceImage(object):
    """A packed image known only by its device address (e.g. a peer's staging
    buffer mapped with ipc_open); quacks like the uint8 tensor the restore
    paths expect."""
    is_cuda = True

    def __init__(self, ptr, nbytes):
        self._ptr, self._n = ptr, nbytes

    def data_ptr(self):
        return self._ptr

    def numel(self):
        return self._n


def enable_peer(device, peer):
    check(lib().fcp_enable_peer(device, peer))


def host_unregister(addr):
    check(lib().fcp_host_unregister(addr))


def chunk_xxh64(packed_ptr, total_bytes, out_ptr, chunk_bytes=DEFAULT_CHUNK_BYTES, stream=0):
    """XXH64 of every chunk of a device-resident image into device memory."""
    check(lib().fcp_chunk_xxh64(packed_ptr, total_bytes, chunk_bytes, out_ptr, stream))


def host_chunk_crc32c(buf, total_bytes=None, chunk_bytes=DEFAULT_CHUNK_BYTES, threads=0):
    """Standard CRC-32C of every chunk of a packed image in host memory (a
    buffer object or an address); host code only, no CUDA."""
    if isinstance(buf, int):
        addr = buf
    else:
        view = np.frombuffer(buf, dtype=np.uint8)
        addr = view.ctypes.data
        if total_bytes is None:
            total_bytes = view.size
    n = (total_bytes + chunk_bytes - 1) // chunk_bytes
    out = np.zeros(n, dtype=np.uint32)
    rc = lib().fcp_host_chunk_crc32c(addr, total_bytes, chunk_bytes, out.ctypes.data, threads)
    if rc:
        raise FcpError(rc, "fcp_host_chunk_crc32c failed")
    return out


def finish_checksums(raw, total_bytes, chunk_bytes=DEFAULT_CHUNK_BYTES):
    raw = np.ascontiguousarray(raw, dtype=np.uint32)
    out = np.zeros_like(raw)
    check(lib().fcp_finish_checksums(raw.ctypes.data, len(raw), total_bytes,
                                     chunk_bytes, out.ctypes.data))
    return out
