"""GPU: kernel-side robustness added in round 2, through the C ABI.

* the L2 bulk prefetch never asks for bytes beyond the end of a tensor or of
  the packed image (sources and images placed at the very end of a cudaMalloc
  allocation);
* pointer refreshes are uploaded on the launching stream, so they are ordered
  before the kernels on a non-blocking stream too;
* the host CRC-32C routine agrees with the sums the pack kernel records.
"""
import numpy as np
import pytest
import torch

from dlrover_wuqiong_b200 import _native
from oracle import oracle

pytestmark = pytest.mark.gpu
CHUNK = 65536


def _raw_alloc(nbytes):
    """A cudaMalloc'ed range outside torch's caching allocator."""
    rt = pytest.importorskip("cuda.bindings.runtime")
    err, ptr = rt.cudaMalloc(nbytes)
    assert int(err) == 0, err
    return rt, int(ptr)


def _as_tensor(ptr, nbytes):
    class _Holder(object):
        __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1",
                                    "data": (ptr, False), "version": 2}
    return torch.as_tensor(_Holder(), device="cuda")


@pytest.mark.parametrize("mis", [0, 4, 13])
def test_sources_and_image_at_the_end_of_an_allocation(mis):
    """Whole-region fast paths with the read side ending exactly where the
    allocation ends: pack reads a tensor that ends there, scatter reads a
    packed image that ends there."""
    torch.cuda.init()
    size = 4 << 20            # a multiple of the 2 MiB mapping granule
    n = 3 * 16384 + 16384     # the tensor: four whole regions
    rt, src = _raw_alloc(size)
    rt2, img = _raw_alloc(size)
    try:
        rng = np.random.default_rng(mis)
        payload = rng.integers(0, 256, n, dtype=np.uint8)
        whole_src = _as_tensor(src, size)
        whole_src.zero_()
        # source [size - n, size); image offset `mis` makes source and image
        # misaligned relative to each other (the realigning paths)
        whole_src[size - n:] = torch.from_numpy(payload).cuda()
        total = mis + n
        pad = (total + 16383) // 16384 * 16384
        whole_img = _as_tensor(img, size)
        whole_img.fill_(0x5A)
        packed_ptr = img + size - pad  # the image's padded end == allocation end
        d = _native.make_descs([src + size - n], [mis], [n])
        plan = _native.Plan(d, total, 0, CHUNK)
        stream = torch.cuda.current_stream().cuda_stream
        plan.pack(packed_ptr, _native.SUM_CRC32C, stream)
        sums = plan.read_checksums(stream)
        expect = np.zeros(total, dtype=np.uint8)
        expect[mis:] = payload
        got = whole_img[size - pad:size - pad + total].cpu().numpy()
        assert got.tobytes() == expect.tobytes()
        assert sums.tolist() == oracle.chunk_crc32c(expect, CHUNK).tolist()
        # scatter: image of exactly `total` bytes ending at the allocation end
        if total % 32 == 0:  # keep the image 32-byte aligned for the kernel's loads
            exact = img + size - total
            whole_img[size - total:] = torch.from_numpy(expect).cuda()
        else:
            exact = packed_ptr
        out = torch.zeros(n, dtype=torch.uint8, device="cuda")
        plan.update_ptrs([out.data_ptr()])
        plan.scatter(exact, _native.SUM_CRC32C, stream)
        sums2 = plan.read_checksums(stream)
        assert out.cpu().numpy().tobytes() == payload.tobytes()
        assert sums2.tolist() == sums.tolist()
        plan.close()
        torch.cuda.synchronize()
    finally:
        rt.cudaFree(src)
        rt2.cudaFree(img)


def test_pointer_refresh_is_ordered_on_a_non_blocking_stream():
    """update_ptrs followed by a launch on a non-blocking side stream: the
    kernels must see the new pointers (the table travels on that stream)."""
    dev = torch.device("cuda", 0)
    side = torch.cuda.Stream(device=dev)  # torch side streams are non-blocking
    n = 1 << 20
    a = [torch.full((n,), i + 1, dtype=torch.uint8, device=dev) for i in range(8)]
    b = [torch.full((n,), 101 + i, dtype=torch.uint8, device=dev) for i in range(8)]
    offs = [i * n for i in range(8)]
    plan = _native.Plan(_native.make_descs([t.data_ptr() for t in a], offs, [n] * 8),
                        8 * n, 0, CHUNK)
    staging = torch.zeros(plan.staging_bytes, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        for rep in range(20):
            src = b if rep % 2 else a
            plan.update_ptrs([t.data_ptr() for t in src])
            plan.pack(staging.data_ptr(), _native.SUM_NONE, side.cuda_stream)
            side.synchronize()
            first = staging[::n].cpu().tolist()
            assert first == [int(t[0]) for t in src], rep
    plan.close()


def test_host_crc_matches_recorded_sums_of_a_save():
    """Agent-side check: the sums the pack kernel records for an image equal
    what the host routine computes over the bytes that reached host memory."""
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(5)
    tensors, offs, sizes, off = [], [], [], 0
    for _ in range(50):
        k = int(rng.integers(1, 700000))
        tensors.append(torch.from_numpy(rng.integers(0, 256, k, dtype=np.uint8)).to(dev))
        off += int(rng.integers(0, 9))
        offs.append(off)
        sizes.append(k)
        off += k
    plan = _native.Plan(_native.make_descs([t.data_ptr() for t in tensors], offs, sizes),
                        off, 0, _native.DEFAULT_CHUNK_BYTES)
    staging = torch.zeros(plan.staging_bytes, dtype=torch.uint8, device=dev)
    host = torch.zeros(off, dtype=torch.uint8).pin_memory()
    stream = torch.cuda.current_stream().cuda_stream
    plan.save_async(staging.data_ptr(), host.data_ptr(), _native.MODE_STAGED,
                    _native.SUM_CRC32C, stream)
    sums = plan.save_wait()
    got = _native.host_chunk_crc32c(host.data_ptr(), off, _native.DEFAULT_CHUNK_BYTES)
    assert got.tolist() == sums.tolist()
    assert got.tolist() == oracle.chunk_crc32c_fast(host.numpy(), _native.DEFAULT_CHUNK_BYTES).tolist()
    plan.close()


# ------------------------------------------------ stripes and sliced scatter
def _random_case(seed, n_tensors, max_bytes, dev):
    rng = np.random.default_rng(seed)
    tensors, offs, sizes, off = [], [], [], 0
    for _ in range(n_tensors):
        k = int(rng.integers(1, max_bytes))
        tensors.append(torch.from_numpy(rng.integers(0, 256, k, dtype=np.uint8)).to(dev))
        off += int(rng.integers(0, 6))  # holes + misalignment
        offs.append(off)
        sizes.append(k)
        off += k
    return tensors, offs, sizes, off


def _oracle_image(tensors, offs, sizes, total):
    host = [t.cpu().numpy() for t in tensors]
    d = np.zeros(len(host), dtype=oracle.DESC_DTYPE)
    d["ptr"] = [h.ctypes.data for h in host]
    d["dst_off"], d["nbytes"] = offs, sizes
    return oracle.pack_descs(d, total)


@pytest.mark.parametrize("n_writers", [2, 3, 8])
def test_striped_save_equals_single_writer_image(n_writers):
    """n plans over the SAME tensors (the ranks of a replicated group), each
    restricted to its chunk window, write one host image: bytes and merged sums
    equal the oracle's."""
    dev = torch.device("cuda", 0)
    tensors, offs, sizes, total = _random_case(40 + n_writers, 60, 900000, dev)
    image = _oracle_image(tensors, offs, sizes, total)
    want_sums = oracle.chunk_crc32c(image, CHUNK)
    n_chunks = len(want_sums)
    per = -(-n_chunks // n_writers)
    host = torch.full((total,), 0x77, dtype=torch.uint8).pin_memory()
    merged = np.zeros(n_chunks, dtype=np.uint32)
    stream = torch.cuda.current_stream().cuda_stream
    plans = []
    for w in range(n_writers):
        plan = _native.Plan(_native.make_descs([t.data_ptr() for t in tensors], offs, sizes),
                            total, 0, CHUNK)
        staging = torch.full((plan.staging_bytes,), 0xEE, dtype=torch.uint8, device=dev)
        w0, w1 = min(n_chunks, w * per), min(n_chunks, (w + 1) * per)
        plan.set_window(w0, w1)
        plan.save_async(staging.data_ptr(), host.data_ptr(), _native.MODE_STAGED,
                        _native.SUM_CRC32C, stream)
        sums = plan.save_wait()
        merged[w0:w1] = sums[w0:w1]
        # nothing outside the window moved
        if w0 > 0:
            assert int(staging[0]) == 0xEE
        plans.append((plan, staging, w0, w1))
    assert host.numpy().tobytes() == image.tobytes()
    assert merged.tolist() == want_sums.tolist()
    # --- sliced scatter: slice s is read from writer s's staging (each holds
    # only its own stripe), one launch, all sums verified
    fresh = [torch.zeros_like(t) for t in tensors]
    reader = _native.Plan(_native.make_descs([t.data_ptr() for t in fresh], offs, sizes),
                          total, 0, CHUNK)
    bases = [st.data_ptr() for _, st, _, _ in plans]
    reader.restore_from_slices(bases, per * CHUNK, _native.SUM_CRC32C, expected=want_sums,
                               read_prefetch=True, stream=stream)
    for a, b in zip(tensors, fresh):
        assert torch.equal(a, b)
    # a wrong byte in one provider's stripe is caught
    _, st1, w0, _ = plans[1]
    st1[w0 * CHUNK + 5] ^= 1
    with pytest.raises(_native.ChecksumMismatch):
        reader.restore_from_slices(bases, per * CHUNK, _native.SUM_CRC32C, expected=want_sums,
                                   read_prefetch=False, stream=stream)
    for plan, _, _, _ in plans:
        plan.close()
    reader.close()


@pytest.mark.parametrize("registered", [False, True])
def test_stage_range_moves_exactly_the_range(registered):
    dev = torch.device("cuda", 0)
    total = (9 << 20) + 12345
    rng = np.random.default_rng(9)
    src = torch.from_numpy(rng.integers(0, 256, total, dtype=np.uint8))
    if registered:
        src = src.pin_memory()
    t = torch.zeros(total, dtype=torch.uint8, device=dev)
    plan = _native.Plan(_native.make_descs([t.data_ptr()], [0], [total]), total, 0, CHUNK)
    staging = torch.full((plan.staging_bytes,), 0xC3, dtype=torch.uint8, device=dev)
    lo, hi = 3 * CHUNK, 3 * CHUNK + (5 << 20) + 77
    plan.stage_range(src.data_ptr(), staging.data_ptr(), lo, hi, registered=registered,
                     stream=torch.cuda.current_stream().cuda_stream)
    got = staging.cpu().numpy()
    assert got[lo:hi].tobytes() == src.numpy()[lo:hi].tobytes()
    assert (got[:lo] == 0xC3).all() and (got[hi:total] == 0xC3).all()
    plan.close()


def test_deferred_copies_of_a_staged_save():
    """fcp_plan_set_defer_copies: pack now, copy when told to (ranks of a big
    node take turns writing into host memory); same bytes, same sums."""
    dev = torch.device("cuda", 0)
    tensors, offs, sizes, total = _random_case(77, 30, 700000, dev)
    image = _oracle_image(tensors, offs, sizes, total)
    plan = _native.Plan(_native.make_descs([t.data_ptr() for t in tensors], offs, sizes),
                        total, 0, CHUNK)
    staging = torch.zeros(plan.staging_bytes, dtype=torch.uint8, device=dev)
    host = torch.full((total,), 0x33, dtype=torch.uint8).pin_memory()
    stream = torch.cuda.current_stream().cuda_stream
    plan.set_defer_copies(True)
    plan.save_async(staging.data_ptr(), host.data_ptr(), _native.MODE_STAGED,
                    _native.SUM_CRC32C, stream)
    plan.save_packed_wait()
    torch.cuda.synchronize()
    assert int(host[0]) == 0x33 and int(host[total - 1]) == 0x33  # nothing copied yet
    with pytest.raises(_native.FcpError, match="deferred"):
        plan.save_wait()
    plan.release_copies()
    sums = plan.save_wait()
    assert host.numpy().tobytes() == image.tobytes()
    assert sums.tolist() == oracle.chunk_crc32c(image, CHUNK).tolist()
    plan.set_defer_copies(False)  # and the undeferred path still works on the same plan
    host.fill_(0)
    plan.save_async(staging.data_ptr(), host.data_ptr(), _native.MODE_STAGED,
                    _native.SUM_CRC32C, stream)
    assert plan.save_wait().tolist() == sums.tolist()
    assert host.numpy().tobytes() == image.tobytes()
    plan.close()
