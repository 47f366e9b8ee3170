"""Raw-image checkpoint files (agent-side persistence fast path).

The reference's saver rebuilds CPU views over the shm and ``torch.save``s them
(dlrover/python/elastic_agent/torch/ckpt_saver.py:914-921): one pickle plus a
zip record (with CRC-32) per tensor.  The shm image is already dense and
carries its own meta tree, so the same information can be written as

    magic "FCPIMG1\\0" | u64 header_len | pickle(header) | pad to 4 KiB | bytes

where ``bytes`` is the byte range of the image that one top-level state name
occupies (contiguous, because offsets run in traversal order, SAVER:255-270)
and ``header["meta"]`` is that name's meta subtree with offsets rebased to the
start of the range.  The reference's FSDP saver already persists its shm raw
(SAVER:1181-1197); this extends the idea to the other savers, opt-in through
``DLROVER_FCP_PERSIST_FORMAT=raw`` in the agent's environment (default ``pt``:
stock ``torch.save`` files, so nothing changes for ``torch.load`` users).

``load_any`` reads either format; ``python -m dlrover_wuqiong_b200.raw_ckpt
SRC DST`` converts a raw file to a ``torch.save`` file.
"""
import mmap
import os
import pickle
import struct
import sys
from concurrent.futures import ThreadPoolExecutor

import torch

from .plan import TensorMeta, traverse_state_dict

MAGIC = b"FCPIMG1\0"
DATA_ALIGN = 4096
PERSIST_FORMAT_ENV = "DLROVER_FCP_PERSIST_FORMAT"
_WRITE_BLOCK = 64 << 20
_WRITERS = 4


def persist_format():
    fmt = os.getenv(PERSIST_FORMAT_ENV, "pt").lower()
    if fmt not in ("pt", "raw"):
        raise ValueError("%s must be 'pt' or 'raw', not %r" % (PERSIST_FORMAT_ENV, fmt))
    return fmt


def subtree_range(meta_subtree):
    """(lo, hi) byte range of the image covered by the non-empty tensor leaves
    below `meta_subtree`; (0, 0) if there are none."""
    span = [None, 0]

    def visit(m):
        if isinstance(m, TensorMeta) and m.numel:
            if span[0] is None or m.offset < span[0]:
                span[0] = m.offset
            span[1] = max(span[1], m.offset + m.numel * m.element_size)
        return m

    traverse_state_dict(meta_subtree, visit)
    return (span[0] or 0, span[1]) if span[0] is not None else (0, 0)


_TENSOR_TAG = "\0fcp_tensor"


def _encode_meta(meta_subtree, base):
    """Meta subtree -> plain containers; a tensor leaf becomes the tuple
    (tag, shape, dtype name, numel, offset - base), so that the file does not
    depend on this package's classes."""
    def visit(m):
        if isinstance(m, TensorMeta):
            return (_TENSOR_TAG, tuple(m.shape), str(m.dtype).split(".", 1)[1], m.numel,
                    m.offset - base if m.numel else 0)
        return m

    return traverse_state_dict(meta_subtree, visit)


def _is_tensor_leaf(v):
    return isinstance(v, tuple) and len(v) == 5 and v[0] == _TENSOR_TAG


class RawImage(object):
    """What the saver hands to ``storage.write_state_dict(..., write_raw)`` in
    place of a state dict: one state name's meta subtree plus the image."""

    def __init__(self, meta_subtree, image_buf, step=0, checksums=None, total=None):
        self.meta = meta_subtree
        self.buf = image_buf
        self.step = step
        self.checksums = checksums  # the image's _FCP_CHECKSUMS meta entry
        self.total = len(image_buf) if total is None else total  # bytes of the whole image

    def __bool__(self):
        return bool(self.meta)


def _round_up(n, a):
    return (n + a - 1) // a * a


def _pwrite_all(fd, view, offset):
    done = 0
    while done < len(view):
        done += os.pwrite(fd, view[done:], offset + done)


def _payload_checksums(raw_image, lo, hi):
    """Checksum entry for payload bytes [lo, hi) of the image: the image's own
    sums when the payload is the whole image, else computed here on the host
    (only images that carry sums get any)."""
    sums = raw_image.checksums
    if not sums or sums.get("algo") != "crc32c" or hi <= lo:
        return None
    if lo == 0 and hi == raw_image.total:
        return sums
    from . import _native
    chunk = int(sums["chunk_bytes"])
    view = memoryview(raw_image.buf)[lo:hi]
    try:
        got = _native.host_chunk_crc32c(view, hi - lo, chunk)
    finally:
        view.release()
    return {"algo": "crc32c", "chunk_bytes": chunk, "sums": got.tolist()}


def write_raw(raw_image, path):
    """Write one state name's byte range and meta to `path` (the
    ``write_func`` of ``CheckpointStorage.write_state_dict``)."""
    lo, hi = subtree_range(raw_image.meta)
    header = {"version": 1, "step": raw_image.step, "nbytes": hi - lo,
              "meta": _encode_meta(raw_image.meta, lo),
              # CRC-32C of the image's fixed windows; window k covers image
              # bytes [k*chunk, (k+1)*chunk), i.e. file payload bytes minus lo
              "image": {"offset": lo, "total": raw_image.total,
                        "checksums": raw_image.checksums},
              # CRC-32C of the windows of THIS file's payload, what a restore
              # from the file verifies
              "payload_checksums": _payload_checksums(raw_image, lo, hi)}
    blob = pickle.dumps(header, protocol=4)
    data_off = _round_up(len(MAGIC) + 8 + len(blob), DATA_ALIGN)
    fd = os.open(path, os.O_WRONLY | os.O_CREAT | os.O_TRUNC, 0o644)
    try:
        _pwrite_all(fd, memoryview(MAGIC + struct.pack("<Q", len(blob)) + blob), 0)
        if hi > lo:
            payload = memoryview(raw_image.buf)[lo:hi]
            blocks = [(o, min(o + _WRITE_BLOCK, hi - lo)) for o in range(0, hi - lo, _WRITE_BLOCK)]
            if len(blocks) == 1:
                _pwrite_all(fd, payload, data_off)
            else:
                # pwrite releases the GIL: a few writers keep the page cache busy
                with ThreadPoolExecutor(_WRITERS) as pool:
                    list(pool.map(lambda b: _pwrite_all(fd, payload[b[0]:b[1]], data_off + b[0]),
                                  blocks))
            payload.release()
        else:
            os.ftruncate(fd, data_off)
    finally:
        os.close(fd)


def is_raw(path):
    try:
        with open(path, "rb") as f:
            return f.read(len(MAGIC)) == MAGIC
    except OSError:
        return False


def read_header(path):
    """-> (header dict, payload offset in the file)"""
    with open(path, "rb") as f:
        head = f.read(len(MAGIC) + 8)
        if head[:len(MAGIC)] != MAGIC:
            raise ValueError("%s is not a raw flash-checkpoint image" % path)
        (n,) = struct.unpack("<Q", head[len(MAGIC):])
        header = pickle.loads(f.read(n))
    if header.get("version") != 1:
        raise ValueError("%s: unknown raw image version %r" % (path, header.get("version")))
    return header, _round_up(len(MAGIC) + 8 + n, DATA_ALIGN)


def read_raw(path):
    """-> the state dict stored in `path`; tensors are private copy-on-write
    views of the file mapping (paged in on first touch, writable)."""
    header, data_off = read_header(path)
    mm = map_payload(path, header, data_off)

    def visit(v):
        if not _is_tensor_leaf(v):
            return v
        _, shape, dtype_name, numel, off = v
        dtype = getattr(torch, dtype_name)
        if numel == 0:
            return torch.tensor([], dtype=dtype)  # as the shm reader does (SAVER:124-125)
        return torch.frombuffer(mm, dtype=dtype, count=numel, offset=off).reshape(shape)

    return traverse_state_dict(header["meta"], visit)


def decode_meta(header):
    """The header's meta tree with TensorMeta leaves (offsets relative to the
    file's payload)."""
    def visit(v):
        if not _is_tensor_leaf(v):
            return v
        _, shape, dtype_name, numel, off = v
        dtype = getattr(torch, dtype_name)
        return TensorMeta(shape=tuple(shape), dtype=dtype, element_size=dtype.itemsize,
                          numel=numel, offset=off)

    return traverse_state_dict(header["meta"], visit)


def map_payload(path, header, data_off):
    """Private writable mapping of the payload bytes (None if there are none)."""
    nbytes = header["nbytes"]
    if not nbytes:
        return None
    if os.path.getsize(path) < data_off + nbytes:
        raise ValueError("%s is truncated: %d payload bytes expected" % (path, nbytes))
    with open(path, "rb") as f:
        return mmap.mmap(f.fileno(), nbytes, access=mmap.ACCESS_COPY, offset=data_off)


def load_any(path, map_location="cpu"):
    """``read_func`` for ``CheckpointStorage.read_state_dict``: raw image or
    ``torch.save`` file, whichever `path` is."""
    if is_raw(path):
        return read_raw(path)
    return torch.load(path, map_location=map_location, weights_only=False)


def convert_to_pt(src, dst):
    """Raw image -> the ``torch.save`` file the reference's saver would have
    written for the same state."""
    torch.save(read_raw(src), dst)


if __name__ == "__main__":
    if len(sys.argv) != 3:
        sys.exit("usage: python -m dlrover_wuqiong_b200.raw_ckpt SRC.raw DST.pt")
    convert_to_pt(sys.argv[1], sys.argv[2])
