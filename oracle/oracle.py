"""TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement of the reference's flash-checkpoint shm layout and copy loop
(paths relative to /root/reference; SAVER =
dlrover/python/elastic_agent/torch/ckpt_saver.py).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / reference arm may
import this module; nothing under ``dlrover_wuqiong_b200/`` does.

Pinned against the reference: ``oracle/gen_golden.py`` runs the reference's own
``SharedMemoryHandler.save_state_dict`` in the build container and commits the
resulting (inputs, meta, shm bytes) under ``tests/golden/``;
``tests/test_oracle.py`` checks this module against those vectors and against
the reference tests' size goldens (9640 B, 1632 B).
"""
import ctypes
import os
import subprocess
from collections import namedtuple
from collections.abc import Mapping

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))

# mirrors the fields of the reference's TensorMeta (SAVER:64-70)
OMeta = namedtuple("OMeta", "shape dtype element_size numel offset")


def traverse(value, visitor):
    """SAVER:94-109: dicts (Mapping) and lists are walked in insertion order;
    everything else — tensors, scalars, *tuples* — is a leaf."""
    if isinstance(value, Mapping):
        return {k: traverse(v, visitor) for k, v in value.items()}
    if isinstance(value, list):
        return [traverse(v, visitor) for v in value]
    return visitor(value)


def layout(state_dict):
    """SAVER:255-270 + SAVER:277-280: offset = running sum of
    numel*element_size in traversal order, no padding.  Returns (meta tree,
    total bytes); non-tensor leaves are carried in the tree unchanged."""
    total = [0]

    def visit(v):
        if not torch.is_tensor(v):
            return v
        m = OMeta(tuple(v.shape), v.dtype, v.element_size(), v.numel(), total[0])
        total[0] += v.numel() * v.element_size()
        return m

    return traverse(state_dict, visit), total[0]


def tensor_bytes(t):
    """Row-major logical contents of a tensor as a uint8 numpy array; this is
    what ``shm_view.copy_(t)`` writes (SAVER:203-206)."""
    t = t.detach().cpu().contiguous()
    if t.numel() == 0:
        return np.zeros(0, dtype=np.uint8)
    return t.reshape(-1).view(torch.uint8).numpy()


def flatten(state_dict):
    """[(tensor, OMeta)] in layout order."""
    meta, total = layout(state_dict)
    out = []

    def walk(v, m):
        if isinstance(v, Mapping):
            for k in v:
                walk(v[k], m[k])
        elif isinstance(v, list):
            for i, x in enumerate(v):
                walk(x, m[i])
        elif torch.is_tensor(v):
            out.append((v, m))

    walk(state_dict, meta)
    return out, meta, total


def pack(state_dict):
    """SAVER:174-206: returns (meta tree, total, shm bytes as uint8 array)."""
    flat, meta, total = flatten(state_dict)
    buf = np.zeros(total, dtype=np.uint8)
    for t, m in flat:
        if m.numel == 0:  # SAVER:201-202
            continue
        n = m.numel * m.element_size
        buf[m.offset:m.offset + n] = tensor_bytes(t)
    return meta, total, buf


def unpack(meta, buf):
    """SAVER:112-137: rebuild the nested dict; tensors are views over ``buf``
    (empty tensors become ``torch.tensor([], dtype)``)."""
    tbuf = torch.from_numpy(buf) if isinstance(buf, np.ndarray) else buf

    def visit(m):
        if not isinstance(m, OMeta):
            return m
        if m.numel == 0:
            return torch.tensor([], dtype=m.dtype)
        n = m.numel * m.element_size
        return tbuf[m.offset:m.offset + n].clone().view(m.dtype).reshape(m.shape)

    return traverse(meta, visit)


# --------------------------------------------------------------- C half
_lib = None


def build_c(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "fcp_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(
            ["gcc", "-O2", "-shared", "-fPIC", "-o", so, src])
    return so


def clib():
    global _lib
    if _lib is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build_c()
        lib = ctypes.CDLL(so)
        lib.oracle_crc32c.restype = ctypes.c_uint32
        lib.oracle_crc32c.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
        lib.oracle_xxh64.restype = ctypes.c_uint64
        lib.oracle_xxh64.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64]
        lib.oracle_chunk_crc32c.restype = None
        lib.oracle_chunk_crc32c.argtypes = [
            ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_void_p]
        lib.oracle_chunk_crc32c_hw.restype = ctypes.c_int
        lib.oracle_chunk_crc32c_hw.argtypes = [
            ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64,
            ctypes.c_uint64, ctypes.c_void_p]
        lib.oracle_pack.restype = None
        lib.oracle_pack.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
        lib.oracle_unpack.restype = None
        lib.oracle_unpack.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
        _lib = lib
    return _lib


def _addr(a):
    return a.ctypes.data if a.size else 0


def crc32c(data):
    a = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
    a = np.ascontiguousarray(a)
    return int(clib().oracle_crc32c(_addr(a), a.size))


def xxh64(data, seed=0):
    a = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
    a = np.ascontiguousarray(a)
    return int(clib().oracle_xxh64(_addr(a), a.size, seed))


def chunk_crc32c(buf, chunk):
    buf = np.ascontiguousarray(buf)
    n = (buf.size + chunk - 1) // chunk
    out = np.zeros(n, dtype=np.uint32)
    if n:
        clib().oracle_chunk_crc32c(_addr(buf), buf.size, chunk, out.ctypes.data)
    return out


def chunk_crc32c_fast(buf, chunk, threads=8):
    """Every chunk's CRC-32C through the crc32 instruction, a few Python
    threads each taking a share of the chunks (ctypes drops the GIL).  Falls
    back to the bytewise routine on a CPU without SSE4.2.  `buf`: any buffer
    object (not copied)."""
    a = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf
    n = (a.size + chunk - 1) // chunk
    out = np.zeros(n, dtype=np.uint32)
    if n == 0:
        return out
    from concurrent.futures import ThreadPoolExecutor
    threads = max(1, min(threads, n))
    per = (n + threads - 1) // threads
    lib = clib()

    def part(t):
        return lib.oracle_chunk_crc32c_hw(_addr(a), a.size, chunk, t * per,
                                          min(n, (t + 1) * per), out.ctypes.data)

    with ThreadPoolExecutor(threads) as pool:
        ok = list(pool.map(part, range(threads)))
    if not all(ok):
        return chunk_crc32c(a, chunk)
    return out


DESC_DTYPE = np.dtype([("ptr", "<u8"), ("dst_off", "<u8"), ("nbytes", "<u8")])


def pack_descs(descs, total):
    """C-level pack of a descriptor table (numpy structured array of
    DESC_DTYPE whose ptr fields are HOST addresses)."""
    out = np.zeros(total, dtype=np.uint8)
    descs = np.ascontiguousarray(descs)
    if len(descs):
        clib().oracle_pack(descs.ctypes.data, len(descs), _addr(out))
    return out


# ------------------------------------------- timed CPU baseline (bench.py only)
def ref_save_into(state_dict, meta, buf):
    """The reference's save loop on host tensors, statement for statement:
    per tensor ``torch.frombuffer(buf, dtype, count, offset).reshape(shape)
    .copy_(t)`` (SAVER:174-206).  `buf` is a writable buffer (mmap)."""
    def walk(v, m):
        if isinstance(v, Mapping):
            for k in v:
                walk(v[k], m[k])
        elif isinstance(v, list):
            for i, x in enumerate(v):
                walk(x, m[i])
        elif torch.is_tensor(v) and v.numel():
            torch.frombuffer(buf, dtype=v.dtype, count=v.numel(),
                             offset=m.offset).reshape(v.shape).copy_(v)

    walk(state_dict, meta)


def ref_load_from(meta, buf, into):
    """The reference's restore: zero-copy views over the buffer
    (SAVER:112-137) followed by the caller's per-tensor ``copy_`` into the
    live tensors (what module/optimizer.load_state_dict do)."""
    def walk(m, t):
        if isinstance(m, Mapping):
            for k in m:
                walk(m[k], t[k])
        elif isinstance(m, list):
            for i, x in enumerate(m):
                walk(x, t[i])
        elif isinstance(m, OMeta) and m.numel:
            view = torch.frombuffer(buf, dtype=m.dtype, count=m.numel,
                                    offset=m.offset).reshape(m.shape)
            t.copy_(view)

    walk(meta, into)
