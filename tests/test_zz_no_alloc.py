"""GPU: fcp_launch_noalloc — the entry point for callers that cannot allocate
device memory (a trainer whose HBM is entirely owned by a caching allocator or
capped per process, an out-of-memory / breakpoint handler with nothing left to
ask for) — against the existing fcp_launch, which takes its per-region scratch
from the caller.

Both go through the C ABI on the same descriptor table.  The outputs are byte
and integer work, so the comparison is exact, as in the other GPU tests.  Two
calls are made back to back with different image sizes: the library's fixed
scratch array keeps the previous call's contents, and the second result must
not depend on them.  This file sorts last on purpose: a run that stops at the
first failure has then already run every other test.
"""
import errno

import numpy as np
import pytest
import torch

from dlrover_wuqiong_b200 import _native
from oracle import oracle

pytestmark = pytest.mark.gpu
CHUNK = 65536


def _case(seed, n_tensors, max_elems):
    """Random odd-sized tensors with a 4-byte hole before each (misaligned
    offsets), descriptor table uploaded to the device."""
    rng = np.random.default_rng(seed)
    dev = torch.device("cuda", 0)
    tensors, offs, sizes, off = [], [], [], 0
    for _ in range(n_tensors):
        n = int(rng.integers(1, max_elems))
        t = torch.from_numpy(rng.integers(0, 256, n, dtype=np.uint8)).to(dev)
        off += 4
        tensors.append(t)
        offs.append(off)
        sizes.append(n)
        off += n
    descs = _native.make_descs([t.data_ptr() for t in tensors], offs, sizes)
    descs_dev = torch.from_numpy(descs.view(np.uint8).copy()).to(dev)
    return tensors, descs_dev, len(descs), off


def _oracle_image(tensors, descs_dev, total):
    """What the reference's copy loop leaves in the buffer for this descriptor
    table (oracle, on host copies of the tensors) and its per-chunk CRC-32C."""
    host = [t.cpu().numpy() for t in tensors]
    d = np.frombuffer(descs_dev.cpu().numpy().tobytes(), dtype=_native.DESC_DTYPE).copy()
    d["ptr"] = [h.ctypes.data for h in host]
    image = oracle.pack_descs(d.view(oracle.DESC_DTYPE), total)
    return image, oracle.chunk_crc32c(image, CHUNK)


def _run(entry, descs_dev, n, total, direction, packed):
    """entry: 'alloc' (existing fcp_launch, caller scratch) or 'noalloc'."""
    lib = _native.lib()
    n_regions = (total + _native.REGION_BYTES - 1) // _native.REGION_BYTES
    n_chunks = (total + CHUNK - 1) // CHUNK
    sums = torch.full((n_chunks,), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    if entry == "alloc":
        scratch = torch.zeros(n_regions, dtype=torch.int32, device="cuda")
        _native.check(lib.fcp_launch(descs_dev.data_ptr(), n, total, CHUNK, 0, n_regions,
                                     packed.data_ptr(), scratch.data_ptr(), sums.data_ptr(),
                                     direction, _native.SUM_CRC32C, stream))
    else:
        _native.check(lib.fcp_launch_noalloc(descs_dev.data_ptr(), n, total, CHUNK,
                                             packed.data_ptr(), sums.data_ptr(), direction,
                                             _native.SUM_CRC32C, stream))
    torch.cuda.synchronize()
    raw = sums.cpu().numpy().view(np.uint32)
    return _native.finish_checksums(raw, total, CHUNK)


def test_noalloc_matches_existing_entry_point_twice_in_a_row():
    # first a large image (fills many scratch slots), then a small one
    for seed, n_tensors, max_elems in ((1, 300, 800000), (2, 40, 9000)):
        tensors, descs_dev, n, total = _case(seed, n_tensors, max_elems)
        assert total <= _native.NOALLOC_MAX_BYTES
        pad = (total + _native.REGION_BYTES - 1) // _native.REGION_BYTES * _native.REGION_BYTES
        packed_a = torch.full((pad,), 0x11, dtype=torch.uint8, device="cuda")
        packed_b = torch.full((pad,), 0x22, dtype=torch.uint8, device="cuda")
        sums_a = _run("alloc", descs_dev, n, total, 0, packed_a)
        sums_b = _run("noalloc", descs_dev, n, total, 0, packed_b)
        assert torch.equal(packed_a[:total], packed_b[:total])
        assert sums_a.tolist() == sums_b.tolist()
        # and both against the oracle, so this is not the CUDA path checked
        # against itself
        want_image, want_sums = _oracle_image(tensors, descs_dev, total)
        assert packed_b[:total].cpu().numpy().tobytes() == want_image.tobytes()
        assert sums_b.tolist() == want_sums.tolist()
        # scatter direction: unpack image b into fresh tensors
        fresh = [torch.zeros_like(t) for t in tensors]
        d = np.frombuffer(descs_dev.cpu().numpy().tobytes(), dtype=_native.DESC_DTYPE).copy()
        d["ptr"] = [t.data_ptr() for t in fresh]
        fresh_descs = torch.from_numpy(d.view(np.uint8).copy()).cuda()
        sums_c = _run("noalloc", fresh_descs, n, total, 1, packed_b)
        assert sums_c.tolist() == sums_a.tolist()
        for a, b in zip(tensors, fresh):
            assert torch.equal(a, b)


def test_noalloc_rejects_images_beyond_its_fixed_scratch():
    lib = _native.lib()
    too_big = _native.NOALLOC_MAX_BYTES + 1
    rc = lib.fcp_launch_noalloc(None, 0, too_big, CHUNK, None, None, 0, _native.SUM_NONE, None)
    assert rc == -errno.E2BIG
    assert b"exceeds the fixed scratch" in lib.fcp_last_error()
