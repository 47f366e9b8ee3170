"""GPU parity: the CUDA pack / scatter / checksum path, called through the
C ABI, against the oracle on the same inputs.  Bit-exact (byte work)."""
import ctypes
import mmap

import numpy as np
import pytest
import torch

from dlrover_wuqiong_b200 import _native
from oracle import oracle

pytestmark = pytest.mark.gpu
CHUNK = _native.DEFAULT_CHUNK_BYTES


def _dev():
    assert torch.cuda.is_available(), "GPU test on a box without CUDA"
    return torch.device("cuda", 0)


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _to_dev(sd):
    return oracle.traverse(sd, lambda v: v.to(_dev()) if torch.is_tensor(v) else v)


def _plan_for(flat, total, chunk=CHUNK):
    """flat: [(device tensor (contiguous), OMeta)]"""
    keep = [t for t, m in flat if m.numel > 0]
    metas = [m for t, m in flat if m.numel > 0]
    d = _native.make_descs([t.data_ptr() for t in keep], [m.offset for m in metas],
                           [m.numel * m.element_size for m in metas])
    return _native.Plan(d, total, 0, chunk), keep


def _pack_and_check(flat, total, expect, chunk=CHUNK, checksum=True):
    plan, keep = _plan_for(flat, total, chunk)
    staging = torch.full((plan.staging_bytes,), 0xA5, dtype=torch.uint8, device=_dev())
    plan.pack(staging.data_ptr(), _native.SUM_CRC32C if checksum else _native.SUM_NONE, _stream())
    torch.cuda.synchronize()
    got = staging[:total].cpu().numpy()
    assert got.tobytes() == expect.tobytes()
    if checksum:
        sums = plan.read_checksums(_stream())
        want = oracle.chunk_crc32c(expect, chunk)
        assert sums.tolist() == want.tolist()
    return plan, staging, keep


def _scatter_and_check(plan, staging, flat, expect_sums=None):
    """Unpack into fresh tensors and compare with the originals."""
    fresh = [torch.zeros_like(t) for t, m in flat if m.numel > 0]
    plan.update_ptrs([t.data_ptr() for t in fresh])
    plan.scatter(staging.data_ptr(), _native.SUM_CRC32C, _stream())
    torch.cuda.synchronize()
    orig = [t for t, m in flat if m.numel > 0]
    for a, b in zip(orig, fresh):
        assert torch.equal(a.view(torch.uint8) if a.dim() else a.reshape(1).view(torch.uint8),
                           b.view(torch.uint8) if b.dim() else b.reshape(1).view(torch.uint8))
    if expect_sums is not None:
        assert plan.read_checksums(_stream()).tolist() == list(expect_sums)


@pytest.mark.parametrize("name", ["simplenet_sgd_9640", "toymodel_adam_1632", "quirks",
                                  "mlp_adamw_stepped", "noncontig"])
def test_pack_matches_reference_golden(golden, name):
    case = golden["cases"][name]
    sd = _to_dev(case["state_dict"])
    flat, meta, total = oracle.flatten(sd)
    flat = [(t.contiguous(), m) for t, m in flat]
    expect = np.frombuffer(case["shm"], dtype=np.uint8)
    plan, staging, keep = _pack_and_check(flat, total, expect, chunk=16384)
    _scatter_and_check(plan, staging, flat, oracle.chunk_crc32c(expect, 16384).tolist())


def _random_case(rng, n, max_elems, dtypes, gap_every=0, misalign_src=False):
    """n tensors of random odd sizes; optional gaps in the layout and sources
    that start at odd byte addresses inside a larger allocation."""
    flat, off = [], 0
    host = []
    for i in range(n):
        dt = dtypes[int(rng.integers(len(dtypes)))]
        numel = int(rng.integers(1, max_elems))
        es = torch.empty((), dtype=dt).element_size()
        nbytes = numel * es
        raw = torch.from_numpy(rng.integers(0, 256, nbytes, dtype=np.uint8))
        if misalign_src:
            pad = int(rng.integers(0, 16 // es)) * es if es < 16 else 0
            big = torch.zeros(nbytes + 64, dtype=torch.uint8, device=_dev())
            big[pad:pad + nbytes] = raw.to(_dev())
            t = big[pad:pad + nbytes].view(dt)
        else:
            t = raw.to(_dev()).view(dt)
        if gap_every and i % gap_every == gap_every - 1:
            off += int(rng.integers(1, 40))
        flat.append((t, oracle.OMeta((numel,), dt, es, numel, off)))
        host.append((raw.numpy(), off))
        off += nbytes
    total = off + (int(rng.integers(0, 20)) if gap_every else 0)
    expect = np.zeros(total, dtype=np.uint8)
    for raw, o in host:
        expect[o:o + raw.size] = raw
    return flat, total, expect


DT = [torch.float32, torch.bfloat16, torch.uint8, torch.float64, torch.int16]


@pytest.mark.parametrize("seed,n,max_elems,gap,mis", [
    (0, 1, 5, 0, False), (1, 50, 9, 0, False), (2, 300, 700, 0, False),
    (3, 300, 700, 7, True), (4, 2000, 40, 3, True), (5, 64, 70000, 0, True),
    (6, 10000, 3, 0, False),
])
def test_pack_scatter_random(seed, n, max_elems, gap, mis):
    rng = np.random.default_rng(seed)
    flat, total, expect = _random_case(rng, n, max_elems, DT, gap, mis)
    plan, staging, _ = _pack_and_check(flat, total, expect, chunk=65536)
    _scatter_and_check(plan, staging, flat, oracle.chunk_crc32c(expect, 65536).tolist())


def test_adamw_like_large_misaligned():
    """4-byte scalar before every big tensor: every big tensor lands at an
    offset that is 4 mod 16 etc.; large enough for the 1024-thread shape."""
    torch.manual_seed(5)
    flat, off = [], 0
    for i in range(12):
        s = torch.full((), float(i), device=_dev())
        flat.append((s, oracle.OMeta((), torch.float32, 4, 1, off)))
        off += 4
        numel = (3 << 20) + 17 * i + 1
        dt = torch.bfloat16 if i % 3 == 2 else torch.float32
        t = torch.randn(numel, device=_dev()).to(dt)
        flat.append((t, oracle.OMeta((numel,), dt, t.element_size(), numel, off)))
        off += numel * t.element_size()
    total = off
    assert total > 148 * 32 * 16384
    expect = np.zeros(total, dtype=np.uint8)
    for t, m in flat:
        expect[m.offset:m.offset + m.numel * m.element_size] = oracle.tensor_bytes(t)
    plan, staging, _ = _pack_and_check(flat, total, expect)
    _scatter_and_check(plan, staging, flat, oracle.chunk_crc32c(expect, CHUNK).tolist())


def test_no_checksum_variant_copies_the_same_bytes():
    rng = np.random.default_rng(9)
    flat, total, expect = _random_case(rng, 40, 90000, DT, 0, True)
    _pack_and_check(flat, total, expect, checksum=False)


def _host_buffer(nbytes):
    """page-aligned anonymous shared mapping, pre-faulted (stands in for the
    /dev/shm segment)"""
    mm = mmap.mmap(-1, max(nbytes, 1), flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)
    arr = np.frombuffer(mm, dtype=np.uint8)
    return mm, arr


@pytest.mark.parametrize("mode", [_native.MODE_STAGED, _native.MODE_DIRECT,
                                  _native.MODE_STAGED_KERNEL])
def test_save_restore_through_host_memory(mode):
    rng = np.random.default_rng(21)
    flat, total, expect = _random_case(rng, 200, 60000, DT, 0, True)
    plan, keep = _plan_for(flat, total)
    plan.set_slice_bytes(4 * CHUNK)  # several slices
    mm, harr = _host_buffer(total)
    harr[:] = 0x5A
    _native.host_register(harr.ctypes.data, total)
    try:
        staging = torch.zeros(plan.staging_bytes, dtype=torch.uint8, device=_dev())
        plan.save_async(staging.data_ptr(), harr.ctypes.data, mode, _native.SUM_CRC32C, _stream())
        plan.save_packed_wait()
        sums = plan.save_wait()
        assert harr[:total].tobytes() == expect.tobytes()
        want = oracle.chunk_crc32c(expect, CHUNK)
        assert sums.tolist() == want.tolist()
        # restore into fresh tensors, verifying the sums
        fresh = [torch.zeros_like(t) for t, m in flat]
        plan.update_ptrs([t.data_ptr() for t in fresh])
        staging.zero_()
        plan.restore(harr.ctypes.data, staging.data_ptr(), mode, _native.SUM_CRC32C,
                     expected=sums, stream=_stream())
        for (a, _), b in zip(flat, fresh):
            assert torch.equal(a.view(torch.uint8), b.view(torch.uint8))
        # corrupt one byte in host memory: the restore must notice
        harr[total // 2] ^= 0x40
        with pytest.raises(_native.ChecksumMismatch):
            plan.restore(harr.ctypes.data, staging.data_ptr(), mode, _native.SUM_CRC32C,
                         expected=sums, stream=_stream())
    finally:
        torch.cuda.synchronize()
        _native.host_unregister(harr.ctypes.data)
        plan.close()
        del harr
        mm.close()


def test_restore_from_device_image():
    rng = np.random.default_rng(33)
    flat, total, expect = _random_case(rng, 100, 50000, DT, 0, False)
    plan, staging, _ = _pack_and_check(flat, total, expect)
    sums = oracle.chunk_crc32c(expect, CHUNK)
    fresh = [torch.zeros_like(t) for t, m in flat]
    plan.update_ptrs([t.data_ptr() for t in fresh])
    plan.restore_from_device(staging.data_ptr(), _native.SUM_CRC32C, expected=sums,
                             stream=_stream())
    for (a, _), b in zip(flat, fresh):
        assert torch.equal(a.view(torch.uint8), b.view(torch.uint8))


def test_empty_plan_and_bad_descriptors():
    plan = _native.Plan(np.zeros(0, dtype=_native.DESC_DTYPE), 0, 0)
    assert plan.num_chunks == 0
    plan.pack(None, _native.SUM_CRC32C, _stream())
    torch.cuda.synchronize()
    bad = _native.make_descs([1, 2], [10, 5], [10, 10])  # overlapping / unordered
    with pytest.raises(_native.FcpError):
        _native.Plan(bad, 100, 0)


@pytest.mark.parametrize("total,chunk", [(1, 64), (31, 64), (32, 64), (33, 64), (64, 64),
                                         (1000, 256), (65536 * 3 + 77, 65536),
                                         (5 * CHUNK + 123457, CHUNK)])
def test_chunk_xxh64_matches_oracle(total, chunk):
    rng = np.random.default_rng(total)
    host = rng.integers(0, 256, total, dtype=np.uint8)
    dev = torch.from_numpy(host).to(_dev())
    n = (total + chunk - 1) // chunk
    out = torch.zeros(n, dtype=torch.int64, device=_dev())
    _native.chunk_xxh64(dev.data_ptr(), total, out.data_ptr(), chunk, _stream())
    torch.cuda.synchronize()
    got = [int(x) & 0xFFFFFFFFFFFFFFFF for x in out.cpu().tolist()]
    want = [oracle.xxh64(host[i * chunk:(i + 1) * chunk]) for i in range(n)]
    assert got == want


def test_image_beyond_4gib_by_properties():
    """Full-size behaviour without a CPU pass over every byte: an image that
    crosses 2^32 (64-bit offsets in descriptors, regions and chunks), checked
    by (1) device-side equality of every tensor's slot in the image, (2) the
    oracle's CRC on sampled chunks (first, last, around 2^32, around tensor
    boundaries), (3) scatter(pack(x)) == x with the same sums (idempotence),
    (4) CRC linearity: crc(a^b) = crc(a)^crc(b)^crc(0) per chunk."""
    torch.manual_seed(77)
    dev = _dev()
    sizes = [(1 << 32) // 4 - 3, 1, (300 << 20) // 4 + 5, 7, (1 << 30) // 4 + 1]  # fp32 elements
    flat, off = [], 0
    for i, n in enumerate(sizes):
        t = torch.empty(n, dtype=torch.float32, device=dev)
        t.view(torch.int32).random_(-2**31, 2**31 - 1)
        flat.append((t, oracle.OMeta((n,), torch.float32, 4, n, off)))
        off += 4 * n
        if i == 1:  # a bf16 scalar: everything behind it sits at 2 mod 4
            s = torch.randn((), device=dev).to(torch.bfloat16)
            flat.append((s, oracle.OMeta((), torch.bfloat16, 2, 1, off)))
            off += 2
    total = off
    assert total > (1 << 32) + (1 << 30)
    plan, keep = _plan_for(flat, total)
    staging = torch.zeros(plan.staging_bytes, dtype=torch.uint8, device=dev)
    plan.pack(staging.data_ptr(), _native.SUM_CRC32C, _stream())
    torch.cuda.synchronize()
    sums = plan.read_checksums(_stream())
    # (1) every tensor's bytes sit at its offset
    for t, m in flat:
        n = m.numel * m.element_size
        assert torch.equal(staging[m.offset:m.offset + n], t.reshape(-1).view(torch.uint8)), m.offset
    # (2) sampled chunks against the oracle
    n_chunks = (total + CHUNK - 1) // CHUNK
    assert len(sums) == n_chunks
    picks = {0, 1, n_chunks - 2, n_chunks - 1, (1 << 32) // CHUNK - 1, (1 << 32) // CHUNK}
    for t, m in flat:
        picks.add(m.offset // CHUNK)
    for c in sorted(picks):
        lo, hi = c * CHUNK, min((c + 1) * CHUNK, total)
        want = oracle.chunk_crc32c(staging[lo:hi].cpu().numpy(), CHUNK)
        assert int(sums[c]) == int(want[0]), c
    # (3) scatter into fresh tensors: same bytes, same sums
    fresh = [torch.zeros_like(t) for t, m in flat]
    plan.update_ptrs([t.data_ptr() for t in fresh])
    plan.scatter(staging.data_ptr(), _native.SUM_CRC32C, _stream())
    torch.cuda.synchronize()
    for (a, _), b in zip(flat, fresh):
        assert torch.equal(a.reshape(-1).view(torch.uint8), b.reshape(-1).view(torch.uint8))
    assert plan.read_checksums(_stream()).tolist() == sums.tolist()
    del fresh, staging
    plan.close()

    # (4) linearity on a 96 MiB image (three packs)
    n = (96 << 20) // 4 + 3
    a = torch.empty(n, dtype=torch.int32, device=dev).random_(-2**31, 2**31 - 1)
    b = torch.empty(n, dtype=torch.int32, device=dev).random_(-2**31, 2**31 - 1)
    crcs = []
    for x in (a, b, a ^ b, torch.zeros_like(a)):
        p, _ = _plan_for([(x, oracle.OMeta((n,), torch.int32, 4, n, 0))], 4 * n)
        st = torch.empty(p.staging_bytes, dtype=torch.uint8, device=dev)
        p.pack(st.data_ptr(), _native.SUM_CRC32C, _stream())
        torch.cuda.synchronize()
        crcs.append(np.asarray(p.read_checksums(_stream()), dtype=np.uint32))
        p.close()
    assert np.array_equal(crcs[2], crcs[0] ^ crcs[1] ^ crcs[3])


@pytest.mark.parametrize("seed,n,max_elems", [(11, 40, 5000), (12, 300, 900), (13, 6, 3_000_000),
                                              (14, 1500, 60), (15, 14, 12_000_000)])
def test_guard_bytes_around_image_and_tensors_stay_intact(seed, n, max_elems):
    """Own bounds check (compute-sanitizer is not available on this pool):
    the image sits between two guard pages and every scatter target between
    two 64-byte guards inside a larger buffer, at odd addresses; after pack
    and scatter (both kernel shapes, realigning paths included) no guard byte
    may have changed and nothing but the targets may have been written."""
    rng = np.random.default_rng(seed)
    dev = _dev()
    G = 64
    flat, bufs, off = [], [], 0
    for i in range(n):
        dt = DT[int(rng.integers(len(DT)))]
        es = torch.empty((), dtype=dt).element_size()
        numel = int(rng.integers(1, max_elems))
        nbytes = numel * es
        lead = G + int(rng.integers(0, 16 // es if es < 16 else 1)) * es  # odd but dtype-aligned
        buf = torch.full((lead + nbytes + G + 16,), 0xEE, dtype=torch.uint8, device=dev)
        buf[lead:lead + nbytes] = torch.from_numpy(
            rng.integers(0, 256, nbytes, dtype=np.uint8)).to(dev)
        t = buf[lead:lead + nbytes].view(dt)
        if i % 3 == 0:
            off += int(rng.integers(1, 9))  # a hole in the image: must stay untouched too
        flat.append((t, oracle.OMeta((numel,), dt, es, numel, off)))
        bufs.append((buf, lead, nbytes))
        off += nbytes
    total = off
    plan, keep = _plan_for(flat, total, chunk=65536)
    if seed == 15:  # large enough for the 1024-thread shape with replicated tables
        assert total > 148 * 32 * 16384, total
    page = 4096
    arena = torch.full((plan.staging_bytes + 2 * page,), 0x77, dtype=torch.uint8, device=dev)
    image = arena[page:page + plan.staging_bytes]
    plan.pack(image.data_ptr(), _native.SUM_CRC32C, _stream())
    torch.cuda.synchronize()
    assert bool((arena[:page] == 0x77).all()) and bool((arena[page + plan.staging_bytes:] == 0x77).all())
    sums = plan.read_checksums(_stream())
    # sources untouched by the pack (guards and payload)
    snap = [b.clone() for b, _, _ in bufs]
    # scatter into fresh guarded targets at the same odd addresses
    targets = []
    for (buf, lead, nbytes), (t, m) in zip(bufs, flat):
        tb = torch.full_like(buf, 0xEE)
        targets.append((tb, tb[lead:lead + nbytes].view(t.dtype)))
    plan.update_ptrs([v.data_ptr() for _, v in targets])
    plan.scatter(image.data_ptr(), _native.SUM_CRC32C, _stream())
    torch.cuda.synchronize()
    assert plan.read_checksums(_stream()).tolist() == sums.tolist()
    assert bool((arena[:page] == 0x77).all()) and bool((arena[page + plan.staging_bytes:] == 0x77).all())
    for (buf, lead, nbytes), s, (tb, _) in zip(bufs, snap, targets):
        assert torch.equal(buf, s)          # the pack wrote nothing into its sources
        assert torch.equal(tb, buf), lead   # payload restored, guards still 0xEE
    plan.close()
