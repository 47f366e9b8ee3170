"""Device side of one rank's shard: owns the native plan (descriptor table on
the GPU), the HBM staging buffer, the small arena through which host-resident
leaves join the packed image, and the pinned mapping of the shm segment.

This is the layer that replaces the reference's per-tensor
``frombuffer(...).copy_(tensor)`` loop (SAVER:197-206) and the per-tensor H2D
on restore.  There is no CPU fallback here: CUDA tensors in the state dict
require libfcp.so and a CUDA device, or an exception is raised.
"""
import contextlib
import ctypes
import os
import threading
import time

import numpy as np
import torch

from . import _native
from .common.log import default_logger as logger

CHECKSUM_KEY = "_FCP_CHECKSUMS"  # extra top-level meta key (ignored by savers)


def _mode_from_env():
    """FCP_TRANSFER_MODE = auto (default: staged, with the device->host leg
    chosen at run time, see SaveLegTuner) | staged (copy engine) |
    staged_kernel (copy kernel for the device->host leg) | direct (pack kernel
    writes the host segment itself)."""
    return {"direct": _native.MODE_DIRECT,
            "staged_kernel": _native.MODE_STAGED_KERNEL}.get(
                os.getenv("FCP_TRANSFER_MODE", "auto"), _native.MODE_STAGED)


class SaveLegTuner(object):
    """Chooses the device->host leg of staged saves from what this process
    measures on this host.

    The copy engine is free (no SMs) and sustains ~57 GB/s on a Gen5 x16 link,
    so it is the default.  On some multi-GPU hosts concurrent copy-engine
    *writes* into host memory stop scaling (measured: 33.5 GB/s/rank at 2
    GPUs, 9.8 at 8) while SM stores into the mapped segment keep more of the
    link (51 and 11.6 GB/s/rank).  So: if a copy-engine save of >= 1 GiB runs
    below GOOD_GBPS, the next save tries the copy kernel and the faster of the
    two (by > 10 %) is kept.  The very first save of a process is not used as
    a sample."""

    MIN_BYTES = 1 << 30
    GOOD_GBPS = 45.0  # 0.8 x what the copy engine sustains on an idle link

    def __init__(self):
        self.state = "warmup"  # -> engine_probe -> [kernel_probe] -> settled
        self.rates = {}
        self.choice = _native.MODE_STAGED

    @property
    def settled(self):
        return self.state == "settled"

    def mode(self):
        return _native.MODE_STAGED_KERNEL if self.state == "kernel_probe" else self.choice

    def kernel_leg_failed(self):
        self.state, self.choice = "settled", _native.MODE_STAGED

    def report(self, mode, nbytes, seconds):
        if self.settled or nbytes < self.MIN_BYTES or seconds <= 0:
            return
        rate = nbytes / seconds / 1e9
        if self.state == "warmup":
            self.state = "engine_probe"
        elif self.state == "engine_probe":
            self.rates["engine"] = rate
            self.state = "settled" if rate >= self.GOOD_GBPS else "kernel_probe"
        elif self.state == "kernel_probe" and mode == _native.MODE_STAGED_KERNEL:
            self.rates["kernel"] = rate
            if rate > 1.1 * self.rates["engine"]:
                self.choice = _native.MODE_STAGED_KERNEL
            self.state = "settled"
            logger.info("device->host leg: copy engine %.1f GB/s, copy kernel %.1f GB/s -> %s",
                        self.rates["engine"], rate,
                        "copy kernel" if self.choice == _native.MODE_STAGED_KERNEL else "copy engine")


class DmaGate(object):
    """Turn-taking for the device->host copies of the ranks of one node.

    Measured on an 8-GPU host (tools/hostlink_matrix.py,
    profiles/r02_hostlink.md): copy-engine writes into host memory reach
    90 GB/s with all 8 ranks writing at once but 102 GB/s when 4 write at a
    time (2 at a time: 85, one: 57); reads scale to 186 GB/s with all 8.  So
    on nodes with more than `FCP_DMA_WRITERS` (default 4) local ranks, a
    rank's copies wait for one of that many slots: rank r shares slot
    r mod k with the ranks k apart and they alternate.  A slot is an advisory
    lock on a file in /dev/shm; it is only held while the copies run, never
    while the training thread is blocked (the pack kernels run first)."""

    def __init__(self, local_rank, local_world):
        k = int(os.getenv("FCP_DMA_WRITERS", "4"))
        self.enabled = 0 < k < local_world
        self._fd = -1
        self.wait_s = 0.0
        if self.enabled:
            run_id = os.getenv("TORCHELASTIC_RUN_ID", "")
            path = "/dev/shm/%sfcp_dma_slot_%d" % (run_id + "_" if run_id else "", local_rank % k)
            try:
                self._fd = os.open(path, os.O_CREAT | os.O_RDWR, 0o600)
            except OSError as e:
                logger.warning("cannot open %s (%s): device->host copies are not paced", path, e)
                self.enabled = False

    def acquire(self):
        if self.enabled:
            import fcntl
            t0 = time.perf_counter()
            fcntl.flock(self._fd, fcntl.LOCK_EX)
            self.wait_s = time.perf_counter() - t0

    def release(self):
        if self.enabled:
            import fcntl
            fcntl.flock(self._fd, fcntl.LOCK_UN)

    def close(self):
        if self._fd >= 0:
            os.close(self._fd)
            self._fd = -1


def default_bounce_workers():
    """Host threads that feed the pinned ring of an unpinned restore: half the
    cores, shared by the local ranks, between 2 and 8 (8 ranks x 8 threads on
    a 32-core host only fought each other)."""
    env = int(os.getenv("FCP_BOUNCE_WORKERS", "0"))
    if env > 0:
        return env
    cores = os.cpu_count() or 1
    local_world = max(1, int(os.getenv("LOCAL_WORLD_SIZE", "1") or 1))
    return max(2, min(8, cores // (2 * local_world)))


def buffer_address(buf):
    """Address of a writable buffer object (mmap / memoryview)."""
    return ctypes.addressof(ctypes.c_char.from_buffer(buf))


REGISTER_PIECE_BYTES = 64 << 20


def _memmove_all(jobs, workers=4, parallel_above=64 << 20):
    """jobs: [(dst address, src address, nbytes)], disjoint."""
    if sum(j[2] for j in jobs) < parallel_above or len(jobs) < 2:
        for dst, src, n in jobs:
            ctypes.memmove(dst, src, n)
        return
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(workers) as pool:
        list(pool.map(lambda j: ctypes.memmove(j[0], j[1], j[2]), jobs))


class HostMapping(object):
    """cudaHostRegister'ed view of the shm segment.

    Staged mode pins the segment in pieces of REGISTER_PIECE_BYTES: pinning is
    serial inside the driver (~12 GB/s) and blocks other CUDA calls of the
    process while it runs, so a restarted trainer pins in the background, one
    piece at a time with a pause in between, and its training steps only ever
    wait for one piece.  Direct mode needs one registration for the kernel's
    alias and pins everything at once."""

    def __init__(self, buf, nbytes, piece_bytes=REGISTER_PIECE_BYTES, byte_range=None):
        """`byte_range` = (lo, hi): pin only the pieces that overlap it (a rank
        that writes one stripe of a shared image pins that stripe alone)."""
        self.addr = buffer_address(buf)
        self.nbytes = nbytes
        self.piece = piece_bytes
        self._done = []
        self.register_s = 0.0
        self.complete = False
        lo, hi = byte_range if byte_range is not None else (0, nbytes)
        step = self.piece if self.piece else max(self.nbytes, 1)
        self.lo = (lo // step) * step
        self.hi = min(nbytes, -(-hi // step) * step)

    def covers(self, lo, hi):
        return self.complete and self.lo <= lo and hi <= self.hi

    def register_all(self, pause=0.0, ready=None):
        """`ready(upto)`: called before a piece is pinned; returns once the
        first `upto` bytes of the segment are resident (a segment that is still
        being populated in the background, SharedMemory.wait_populated)."""
        t0 = time.time()
        step = self.piece if self.piece else max(self.nbytes, 1)
        for off in range(self.lo, self.hi, step):
            if ready is not None:
                ready(min(off + step, self.nbytes))
            _native.host_register(self.addr + off, min(step, self.nbytes - off))
            self._done.append(off)
            if pause:
                time.sleep(pause)
        self.register_s = time.time() - t0
        self.complete = True
        return self

    def close(self):
        done, self._done = self._done, []
        self.complete = False
        for off in done:
            try:
                _native.host_unregister(self.addr + off)
            except Exception as e:  # noqa: BLE001
                logger.warning("host_unregister failed: %s", e)


class DeviceShardEngine(object):
    """Moves one shard between live tensors and a host buffer."""

    def __init__(self, device, chunk_bytes=_native.DEFAULT_CHUNK_BYTES, checksum=True):
        if not torch.cuda.is_available():
            raise RuntimeError(
                "the state dict holds CUDA tensors but CUDA is not available; "
                "the flash-checkpoint GPU path has no CPU fallback")
        _native.lib()  # raises ImportError if the extension is missing
        self.device = torch.device(device)
        self.chunk_bytes = chunk_bytes
        self.checksum = _native.SUM_CRC32C if checksum else _native.SUM_NONE
        self.mode = _mode_from_env()
        self.tuner = SaveLegTuner() if os.getenv("FCP_TRANSFER_MODE", "auto") == "auto" else None
        # optional second algorithm: standard XXH64 per chunk, computed on the
        # staged image after the save is durable and re-checked after a restore
        self.xxh64 = os.getenv("FCP_XXH64", "0") == "1" and self.mode != _native.MODE_DIRECT
        self._plan = None
        self._layout_key = None
        self._staging = None
        self._arena_dev = None
        self._arena_pin = None
        self._mapping = None
        self._mapping_job = None  # background registration after a cold restore
        self._keepalive = []
        self._gate = DmaGate(int(os.getenv("LOCAL_RANK", "0") or 0),
                             int(os.getenv("LOCAL_WORLD_SIZE", "1") or 1))
        self._save_bytes = 0
        self._window = None
        self._restore_host_leaves = []
        self._peer_maps = {}  # rank -> (ipc handle, mapped base) of peers' staging
        self._warned_host_leaves = False
        self.stats = {}

    # -- plan -------------------------------------------------------------
    def _sources(self, leaves):
        """Device pointers of all non-empty leaves; host leaves are staged
        through the arena, non-contiguous ones through a temporary."""
        self._keepalive = []
        host = [lf for lf in leaves if lf.nbytes and not lf.is_cuda]
        host_bytes = sum(lf.nbytes for lf in host)
        if host_bytes:
            if host_bytes > (256 << 20) and not self._warned_host_leaves:
                self._warned_host_leaves = True
                logger.warning(
                    "%.1f GB of this state dict lives in host memory; it joins the image through "
                    "a pinned arena and the GPU (so that one kernel checksums every byte), which "
                    "costs two extra trips over the host link. Keep large tensors on the GPU "
                    "while checkpointing (e.g. ShardedOptimStateDictConfig(offload_to_cpu=False)).",
                    host_bytes / 1e9)
            if self._arena_pin is None or self._arena_pin.numel() < host_bytes:
                self._arena_pin = torch.empty(host_bytes, dtype=torch.uint8).pin_memory()
                self._arena_dev = torch.empty(host_bytes, dtype=torch.uint8,
                                              device=self.device)
        ptrs, offs, sizes, cursor = [], [], [], 0
        for lf in leaves:
            if not lf.nbytes:
                continue
            t = lf.tensor
            if lf.is_cuda:
                if t.device != self.device:
                    raise RuntimeError("tensor on %s, engine on %s" % (t.device, self.device))
                if not t.is_contiguous():
                    t = t.contiguous()  # row-major contents, as copy_ would write
                    self._keepalive.append(t)
                ptrs.append(t.data_ptr())
            else:
                ptrs.append(self._arena_dev.data_ptr() + cursor)
                cursor += lf.nbytes
            offs.append(lf.meta.offset)
            sizes.append(lf.nbytes)
        return ptrs, offs, sizes, host

    def _stage_host_leaves(self, host, to_device):
        """Host leaves <-> arena (pinned), one async copy for all of them."""
        if not host:
            return
        n = sum(lf.nbytes for lf in host)
        pin = self._arena_pin[:n]
        base = pin.data_ptr()
        # plain memmove per leaf (GIL released), a few at a time when there is
        # a lot to move: torch's copy_ fans every 2 MB copy out over all OpenMP
        # threads, which crawls (~1 GB/s measured) once several ranks and the
        # agent share the cores
        jobs, keep, cursor = [], [], 0
        if to_device:
            for lf in host:
                src = lf.tensor.detach()
                if not src.is_contiguous():
                    src = src.contiguous()
                    keep.append(src)
                jobs.append((base + cursor, src.data_ptr(), lf.nbytes))
                cursor += lf.nbytes
            _memmove_all(jobs)
            self._arena_dev[:n].copy_(pin, non_blocking=True)
        else:
            pin.copy_(self._arena_dev[:n], non_blocking=False)
            strided = []
            for lf in host:
                t = lf.tensor
                if t.is_contiguous():
                    jobs.append((t.data_ptr(), base + cursor, lf.nbytes))
                else:
                    strided.append((t, cursor))
                cursor += lf.nbytes
            _memmove_all(jobs)
            for t, at in strided:
                with torch.no_grad():
                    t.copy_(pin[at:at + t.numel() * t.element_size()].view(t.dtype).reshape(t.shape))

    def _ensure_plan(self, ptrs, offs, sizes, total):
        key = (total, tuple(offs), tuple(sizes))
        if self._plan is None or self._layout_key != key:
            if self._plan is not None:
                self._plan.close()
            descs = _native.make_descs(ptrs, offs, sizes)
            self._plan = _native.Plan(descs, total, self.device.index or 0, self.chunk_bytes)
            self._layout_key = key
            slice_bytes = int(os.getenv("FCP_SLICE_BYTES", "0"))
            if slice_bytes:
                self._plan.set_slice_bytes(slice_bytes)
        else:
            self._plan.update_ptrs(ptrs)
        need = self._plan.staging_bytes
        if self.mode != _native.MODE_DIRECT and (
                self._staging is None or self._staging.numel() < need):
            self._staging = None
            self._staging = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._plan

    def _join_mapping_job(self):
        job = self._mapping_job
        if job is not None:
            job.join()
            self._mapping_job = None

    def _register_in_background(self, buf, total):
        """Pin the segment off the critical path (a restarted trainer's first
        save would otherwise pay ~0.08 s/GB for it)."""
        mapping = HostMapping(buf, total, self._piece_bytes())

        def run():
            try:
                with torch.cuda.device(self.device):
                    mapping.register_all(pause=0.002)
                self._mapping = mapping
                self.stats["register_s"] = mapping.register_s
            except Exception as e:  # noqa: BLE001 - retried synchronously by the next save
                mapping.close()
                logger.warning("background registration of the shm failed: %s", e)

        self._mapping_job = threading.Thread(target=run, name="fcp-register", daemon=True)
        self._mapping_job.start()

    def _piece_bytes(self):
        return 0 if self.mode == _native.MODE_DIRECT else REGISTER_PIECE_BYTES

    def _ensure_mapping(self, buf, total, ready=None, byte_range=None):
        self._join_mapping_job()
        addr = buffer_address(buf)
        lo, hi = byte_range if byte_range is not None else (0, total)
        if self._mapping is not None and (self._mapping.addr != addr
                                          or self._mapping.nbytes < total
                                          or not self._mapping.covers(lo, hi)):
            self._mapping.close()
            self._mapping = None
        if self._mapping is None and total:
            self._mapping = HostMapping(buf, total, self._piece_bytes(),
                                        byte_range).register_all(ready=ready)
            self.stats["register_s"] = self._mapping.register_s
        if self._plan is not None:
            self._plan.set_copy_piece_bytes(self._piece_bytes())
        return addr

    # -- save -------------------------------------------------------------
    @staticmethod
    def _nvtx(name):
        """NVTX range around a phase when FCP_NVTX=1 (for nsys timelines)."""
        if os.getenv("FCP_NVTX", "0") != "1":
            return contextlib.nullcontext()
        return torch.cuda.nvtx.range(name)

    def stripe_of(self, total, index, count):
        """(chunk_begin, chunk_end) of stripe `index` when an image of `total`
        bytes is cut into `count` runs of whole chunks, as equal as possible."""
        n_chunks = -(-total // self.chunk_bytes)
        per = -(-n_chunks // count)
        return min(n_chunks, index * per), min(n_chunks, (index + 1) * per)

    def save_begin(self, leaves, total, buf, ready=None, window=None):
        """Enqueue pack + transfer of the shard into `buf`; returns once the
        work is enqueued.  Call save_packed_wait() before mutating the tensors
        and save_finish() for durability + checksums.  `ready`: see
        HostMapping.register_all.  `window` = (chunk_begin, chunk_end): move
        only that stripe of the image (a replicated group where every rank
        writes one stripe over its own host link); staged modes only."""
        with torch.cuda.device(self.device):
            t0 = time.perf_counter()
            with self._nvtx("fcp.plan"):
                ptrs, offs, sizes, host = self._sources(leaves)
                plan = self._ensure_plan(ptrs, offs, sizes, total)
            byte_range = None
            if window is not None:
                if self.mode == _native.MODE_DIRECT:
                    raise ValueError("striped saves need a staged transfer mode")
                byte_range = (window[0] * self.chunk_bytes,
                              min(total, window[1] * self.chunk_bytes))
                plan.set_window(window[0], window[1])
            else:
                plan.set_window()
            self._window = window
            with self._nvtx("fcp.pin"):
                addr = self._ensure_mapping(buf, total, ready, byte_range)
            stream = torch.cuda.current_stream(self.device)
            with self._nvtx("fcp.host_leaves"):
                self._stage_host_leaves(host, to_device=True)
            t1 = time.perf_counter()
            staging_ptr = self._staging.data_ptr() if self._staging is not None else None
            if self.tuner is not None:
                self.mode = self.tuner.mode()
            # paced copies (DmaGate): pack now, copy when this rank has a slot
            plan.set_defer_copies(self._gate.enabled and self.mode != _native.MODE_DIRECT)
            try:
                plan.save_async(staging_ptr, addr, self.mode, self.checksum, stream.cuda_stream)
            except _native.FcpError:
                if self.tuner is None or self.mode != _native.MODE_STAGED_KERNEL:
                    raise
                # the host pointer is not usable from the device as is: stay
                # with the copy engine (nothing was launched)
                self.tuner.kernel_leg_failed()
                self.mode = _native.MODE_STAGED
                plan.save_async(staging_ptr, addr, self.mode, self.checksum, stream.cuda_stream)
            self._save_bytes = total if byte_range is None else max(0, byte_range[1] - byte_range[0])
            self.stats["plan_s"] = t1 - t0

    def save_packed_wait(self):
        with self._nvtx("fcp.pack_wait"):
            self._plan.save_packed_wait()

    def save_finish(self):
        paced = self._gate.enabled and self.mode != _native.MODE_DIRECT
        if paced:
            self._gate.acquire()
        try:
            if paced:
                self._plan.release_copies()
            with self._nvtx("fcp.d2h_wait"):
                sums = self._plan.save_wait(want_sums=self.checksum != _native.SUM_NONE)
        finally:
            if paced:
                self._gate.release()
        self.stats["dma_slot_wait_s"] = self._gate.wait_s if paced else 0.0
        self._keepalive = []
        self.stats.update(self._plan.timings())
        self.stats["save_mode"] = self.mode
        if self.tuner is not None:
            self.tuner.report(self.mode, self._save_bytes, self.stats.get("total_ms", 0.0) / 1e3)
            if self.tuner.settled:
                self.mode = self.tuner.choice
        self.stats["xxh64"] = self._staged_xxh64() if self.xxh64 else None
        total_ms, kernel_ms = self.stats.get("total_ms", 0.0), self.stats.get("kernel_ms", 0.0)
        if total_ms > 0 and self._save_bytes:
            # the reference logs seconds only (FC/engine.py:94-105); rates tell
            # whether the link or the host is the limit
            self.stats["save_GBps"] = self._save_bytes / total_ms / 1e6
            logger.info("flash checkpoint save: %.2f GB in shm after %.1f ms (%.1f GB/s over the "
                        "host link), tensors free after %.1f ms (pack %.0f GB/s of HBM traffic)",
                        self._save_bytes / 1e9, total_ms, self.stats["save_GBps"], kernel_ms,
                        2 * self._save_bytes / kernel_ms / 1e6 if kernel_ms > 0 else 0.0)
        return sums

    def _staged_xxh64(self):
        """XXH64 (seed 0) of every chunk of the image now in HBM staging."""
        plan = self._plan
        if plan is None or self._staging is None or plan.total_bytes == 0:
            return []
        with torch.cuda.device(self.device):
            out = torch.empty(plan.num_chunks, dtype=torch.int64, device=self.device)
            stream = torch.cuda.current_stream(self.device)
            _native.chunk_xxh64(self._staging.data_ptr(), plan.total_bytes, out.data_ptr(),
                                self.chunk_bytes, stream.cuda_stream)
            return [int(x) & 0xFFFFFFFFFFFFFFFF for x in out.cpu().tolist()]

    # -- restore ----------------------------------------------------------
    def restore(self, leaves, total, buf, expected=None, expected_xxh64=None, transient=False):
        """Fill the tensors of `leaves` from the packed image in `buf`.
        `transient`: `buf` is not the shm segment (e.g. a mapped checkpoint
        file): stream it through the pinned ring, pin nothing."""
        with torch.cuda.device(self.device):
            for lf in leaves:
                if lf.is_cuda and lf.nbytes and not lf.tensor.is_contiguous():
                    raise ValueError("restore target must be contiguous")
            t_begin = time.perf_counter()
            ptrs, offs, sizes, host = self._sources(leaves)
            plan = self._ensure_plan(ptrs, offs, sizes, total)
            stream = torch.cuda.current_stream(self.device)
            if transient and total and self._staging is None:  # direct mode keeps none
                self._staging = torch.empty(plan.staging_bytes, dtype=torch.uint8,
                                            device=self.device)
            staging_ptr = self._staging.data_ptr() if self._staging is not None else None
            checksum = self.checksum if expected is not None else _native.SUM_NONE
            self._join_mapping_job()
            pin_later = False
            t_ready = time.perf_counter()
            cold = (self._mapping is None or self._mapping.addr != buffer_address(buf)
                    or self._mapping.nbytes < total or not self._mapping.covers(0, total))
            if transient:
                if total:
                    plan.restore_bounced(buffer_address(buf), staging_ptr, checksum, expected,
                                         default_bounce_workers(), stream.cuda_stream)
                self.stats["restore_path"] = "bounced_file"
            elif (cold and total and self.mode != _native.MODE_DIRECT
                    and os.getenv("FCP_COLD_RESTORE", "bounce") == "bounce"):
                # fresh process, segment not pinned yet: stream it through the
                # pinned ring now, pin it in the background for later saves
                if self._mapping is not None:
                    self._mapping.close()
                    self._mapping = None
                plan.restore_bounced(buffer_address(buf), staging_ptr, checksum, expected,
                                     default_bounce_workers(), stream.cuda_stream)
                self.stats["restore_path"] = "bounced"
                pin_later = True
            else:
                addr = self._ensure_mapping(buf, total)
                plan.restore(addr, staging_ptr, self.mode, checksum, expected,
                             stream.cuda_stream)
                self.stats["restore_path"] = "registered"
            t_moved = time.perf_counter()
            self._stage_host_leaves(host, to_device=False)
            if pin_later:
                self._register_in_background(buf, total)
            self.stats.update(self._plan.timings())
            self.stats["restore_prepare_s"] = t_ready - t_begin
            self.stats["restore_move_s"] = t_moved - t_ready
            self.stats["restore_host_leaves_s"] = time.perf_counter() - t_moved
            if total:
                wall = time.perf_counter() - t_begin
                self.stats["restore_GBps"] = total / wall / 1e9
                logger.info("flash checkpoint restore (%s): %.2f GB into live tensors in %.1f ms "
                            "(%.1f GB/s)", self.stats["restore_path"], total / 1e9, wall * 1e3,
                            self.stats["restore_GBps"])
            if expected_xxh64 is not None and self.xxh64:
                got = self._staged_xxh64()
                if got != list(expected_xxh64):
                    bad = next(i for i, (a, b) in enumerate(zip(got, expected_xxh64)) if a != b)
                    raise _native.ChecksumMismatch(
                        -74, "XXH64 mismatch in chunk %d: expected %016x, got %016x"
                        % (bad, expected_xxh64[bad], got[bad]))

    def restore_from_device(self, leaves, total, packed, expected=None):
        """Same, but the packed image is a device uint8 tensor (peer recovery)."""
        with torch.cuda.device(self.device):
            for lf in leaves:
                if lf.is_cuda and lf.nbytes and not lf.tensor.is_contiguous():
                    raise ValueError("restore target must be contiguous")
            ptrs, offs, sizes, host = self._sources(leaves)
            plan = self._ensure_plan(ptrs, offs, sizes, total)
            stream = torch.cuda.current_stream(self.device)
            checksum = self.checksum if expected is not None else _native.SUM_NONE
            peer = isinstance(packed, _native.RawDeviceImage)  # a peer's HBM over NVLink
            plan.set_read_prefetch(not peer)
            try:
                plan.restore_from_device(packed.data_ptr(), checksum, expected,
                                         stream.cuda_stream)
            finally:
                plan.set_read_prefetch(True)
            self._stage_host_leaves(host, to_device=False)

    # -- striped restore of a replicated group ------------------------------
    def prepare_restore(self, leaves, total):
        """Plan + staging for a restore whose bytes arrive in pieces
        (stage_range / restore_from_slices).  Returns the staging tensor."""
        with torch.cuda.device(self.device):
            for lf in leaves:
                if lf.is_cuda and lf.nbytes and not lf.tensor.is_contiguous():
                    raise ValueError("restore target must be contiguous")
            ptrs, offs, sizes, host = self._sources(leaves)
            plan = self._ensure_plan(ptrs, offs, sizes, total)
            plan.set_window()
            if total and self._staging is None:
                self._staging = torch.empty(plan.staging_bytes, dtype=torch.uint8,
                                            device=self.device)
            self._restore_host_leaves = host
            return self._staging

    def stage_range(self, buf, total, lo, hi):
        """Host image bytes [lo, hi) -> the same offsets of the HBM staging
        (copy engine if the segment is pinned there, pinned ring otherwise)."""
        with torch.cuda.device(self.device):
            self._join_mapping_job()
            addr = buffer_address(buf)
            pinned = (self._mapping is not None and self._mapping.addr == addr
                      and self._mapping.covers(lo, hi))
            if pinned:
                self._plan.set_copy_piece_bytes(self._piece_bytes())
            stream = torch.cuda.current_stream(self.device)
            self._plan.stage_range(addr, self._staging.data_ptr(), lo, hi, registered=pinned,
                                   n_workers=default_bounce_workers(),
                                   stream=stream.cuda_stream)
            self.stats["stage_path"] = "registered" if pinned else "bounced"
            self.stats["stage_ms"] = self._plan.timings()["total_ms"]

    def restore_from_slices(self, total, bases, slice_bytes, expected=None, local=False,
                            first_slice=0):
        """One scatter launch over an image whose slice s is read through
        bases[s] (own staging or peers' staging mapped with CUDA IPC); call
        prepare_restore first.  `local`: every base is this GPU's memory."""
        with torch.cuda.device(self.device):
            stream = torch.cuda.current_stream(self.device)
            checksum = self.checksum if expected is not None else _native.SUM_NONE
            self._plan.restore_from_slices(bases, slice_bytes, checksum, expected,
                                           read_prefetch=local, first_slice=first_slice,
                                           stream=stream.cuda_stream)
            self._stage_host_leaves(self._restore_host_leaves, to_device=False)
            self._restore_host_leaves = []
            self.stats.update(self._plan.timings())
            self.stats["restore_path"] = "slices_local" if local else "slices_nvlink"

    def peer_base(self, rank, handle):
        """Device address, in this process, of the staging allocation rank
        `rank` exported as `handle` (CUDA IPC).  Mappings are kept for the life
        of the engine and reused: a peer's staging buffer lives as long as its
        layout, while opening and above all closing a mapping of tens of GB
        costs 0.1 s and more (measured: 0.11 s to close 2 GiB) — too much to
        pay on every restore.  One mapping per peer; a changed handle replaces
        it."""
        held = self._peer_maps.get(rank)
        if held is not None and held[0] == handle:
            return held[1]
        if held is not None:
            try:
                _native.ipc_close(held[1])
            except Exception as e:  # noqa: BLE001
                logger.warning("ipc_close failed: %s", e)
        with torch.cuda.device(self.device):
            base = _native.ipc_open(handle)
        self._peer_maps[rank] = (handle, base)
        return base

    def _close_peer_maps(self):
        maps, self._peer_maps = self._peer_maps, {}
        for _, base in maps.values():
            try:
                _native.ipc_close(base)
            except Exception as e:  # noqa: BLE001
                logger.warning("ipc_close failed: %s", e)

    @property
    def staging(self):
        return self._staging

    def close(self):
        try:
            if torch.cuda.is_available():
                torch.cuda.synchronize(self.device)
        except Exception:  # noqa: BLE001
            pass
        self._join_mapping_job()
        self._close_peer_maps()
        if self._plan is not None:
            self._plan.close()
            self._plan = None
        if self._mapping is not None:
            self._mapping.close()
            self._mapping = None
        self._staging = None
        self._arena_dev = None
        self._arena_pin = None
        self._gate.close()


def checksums_to_meta(sums, chunk_bytes, xxh64=None, host_leaves=None):
    entry = {"algo": "crc32c", "chunk_bytes": int(chunk_bytes),
             "sums": np.asarray(sums, dtype=np.uint32).tolist()}
    if xxh64 is not None:
        entry["xxh64"] = list(xxh64)
    if host_leaves:
        entry["host_leaves"] = list(host_leaves)
    return entry


def checksums_from_meta(entry, total, chunk_bytes):
    if not entry or entry.get("algo") != "crc32c" or \
            entry.get("chunk_bytes") != chunk_bytes:
        return None
    sums = np.asarray(entry["sums"], dtype=np.uint32)
    if len(sums) != (total + chunk_bytes - 1) // chunk_bytes:
        return None
    return sums
