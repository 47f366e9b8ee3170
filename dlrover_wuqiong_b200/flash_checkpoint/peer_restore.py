"""Restore of REPLICATED state (a DDP group: every rank holds the same state
dict) in which the checkpoint image crosses the host link once per node and
travels between GPUs over NVLink — and, after a rank was replaced while its
peers lived on, does not touch the host at all.

The reference restores replicated state by letting every local rank read the
whole image from shard 0's shm (full_ckpt_engine.py:145-166: n x B bytes over
one host) and recovers a lost shm by broadcasting its bytes as CPU tensors
over gloo (replica.py:275-332, called from engine.py:349-367 on every load).
Here, on a GPU job, ``CheckpointEngine.get_state_dict_from_memory`` calls
``restore_replicated`` instead:

1. every rank reports what it holds of the newest step S: a complete image (or
   the stripe it wrote in a striped save) still sitting in its HBM staging,
   and/or a complete image in its node's shm;
2. the image is cut into one stripe per rank of the node.  Stripe i is
   provided by rank i if its HBM already holds it; else by a rank whose HBM
   holds the whole image (a survivor); else rank i fetches it from shm into its
   staging (1/n of the image over its own host link, all links in parallel);
3. providers export their staging with CUDA IPC; after one barrier every
   rank runs ONE scatter launch that reads each stripe through the provider's
   pointer (plain per-thread loads over NVLink; TMA multicast loads here read
   only this GPU's memory), verifies the CRC-32C of every chunk against the
   sums recorded at save time and unpacks into the live tensors.

Ordering rule: a rank first makes its own stripe complete in
its staging (kept from the last save, or staged from shm), THEN serves; nobody
reads a peer before the barrier that follows everybody's staging, and nobody
frees or rewrites staging before the closing vote.

Everything here is collective over `group` (default: the main group, NCCL);
every decision is computed from the gathered facts, so all ranks take the same
branch.  Returns None when the situation is not one this path handles (no
checksummed image anywhere, a node that can provide neither HBM nor shm for a
stripe — e.g. a replaced node, which takes the reference's replica path) so
the caller falls back, collectively, to the reference behaviour.
"""
import os
import socket
import time

import torch
import torch.distributed as dist

from .. import _native
from .. import plan as _plan
from ..ckpt_saver import DLROVER_CKPT_CONFIG_KEY, FCP_CHECKSUM_KEY, _strip_private
from ..common.log import default_logger as logger
from ..transfer import checksums_from_meta


def _host_id():
    boot = ""
    try:
        with open("/proc/sys/kernel/random/boot_id") as f:
            boot = f.read().strip()
    except OSError:
        pass
    return socket.gethostname() + "/" + boot


def _total_bytes(meta):
    end = 0

    def visit(m):
        nonlocal end
        if isinstance(m, _plan.TensorMeta):
            end = max(end, m.offset + m.numel * m.element_size)
        return m

    _plan.traverse_state_dict(meta, visit)
    return end


def _usable(meta, step):
    conf = meta.get(DLROVER_CKPT_CONFIG_KEY) if meta else None
    return bool(conf is not None and conf.step == step and step > 0
                and not conf.writing_shm and FCP_CHECKSUM_KEY in meta)


def plan_providers(facts, step, total, chunk_bytes):
    """Who provides which stripe, from the gathered facts alone.

    facts[r] = dict(host, hbm=(step, lo, hi), shm_step).  Returns
    {host: [(lo, hi, provider rank, needs_staging)] in stripe order} or None if
    some stripe has no provider.  Pure function (unit-tested on CPU)."""
    hosts = {}
    for r, f in enumerate(facts):
        hosts.setdefault(f["host"], []).append(r)
    n_chunks = -(-total // chunk_bytes)
    out = {}
    for host, ranks in hosts.items():
        n = len(ranks)
        per = -(-n_chunks // n)
        full = [r for r in ranks
                if facts[r]["hbm"][0] == step and facts[r]["hbm"][1] == 0
                and facts[r]["hbm"][2] >= total]
        stripes = []
        for i, r in enumerate(ranks):
            lo = min(total, i * per * chunk_bytes)
            hi = min(total, (i + 1) * per * chunk_bytes)
            h_step, h_lo, h_hi = facts[r]["hbm"]
            if hi <= lo:
                stripes.append((lo, hi, r, False))
            elif h_step == step and h_lo <= lo and hi <= h_hi:
                stripes.append((lo, hi, r, False))
            elif full:
                stripes.append((lo, hi, full[i % len(full)], False))
            elif facts[r]["shm_step"] == step:
                stripes.append((lo, hi, r, True))
            else:
                return None
        out[host] = stripes
    return out


def restore_replicated(handler, into, device, group=None):
    """Collective.  -> (step, state_dict) or None (not handled, see module
    docstring).  Raises ChecksumMismatch on every rank if any rank's image
    failed verification."""
    me = dist.get_rank(group)
    world = dist.get_world_size(group)
    dev = torch.device("cuda", torch.cuda.current_device())
    t_begin = time.perf_counter()

    # ---- 1. what do I hold?
    local_meta = handler.metadata.get(local=True) or {}
    hbm = handler.staged_image()
    try:
        agent_meta = handler.metadata.get() or {}
    except Exception:  # noqa: BLE001 - no agent on this node
        agent_meta = {}
    conf = agent_meta.get(DLROVER_CKPT_CONFIG_KEY)
    shm_step = conf.step if conf is not None and _usable(agent_meta, conf.step) else 0
    if shm_step:
        if handler.shared_memory is None or handler._need_creation or handler._segment_replaced():
            handler.init_shared_memory(create=False)
        if not handler.shared_memory:
            shm_step = 0
    mine = {"host": _host_id(), "hbm": tuple(hbm), "shm_step": shm_step,
            "meta_steps": sorted({s for s, m in ((shm_step, agent_meta), (hbm[0], local_meta))
                                  if _usable(m, s)})}
    facts = [None] * world
    dist.all_gather_object(facts, mine, group=group)
    step = max(max(f["hbm"][0], f["shm_step"]) for f in facts)
    if step <= 0:
        return None

    # ---- 2. the meta tree of step S on every rank
    if _usable(agent_meta, step):
        meta = agent_meta
    elif _usable(local_meta, step):
        meta = local_meta
    else:
        meta = None
    holders = [r for r, f in enumerate(facts) if step in f["meta_steps"]]
    if not holders:
        return None
    if len(holders) < world:
        box = [meta if me == holders[0] else None]
        src = dist.get_global_rank(group, holders[0]) if group else holders[0]
        dist.broadcast_object_list(box, src=src, group=group)
        if meta is None:
            meta = box[0]
    total = _total_bytes(_strip_private(meta))
    entry = meta[FCP_CHECKSUM_KEY]
    chunk = int(entry["chunk_bytes"])
    providers = plan_providers(facts, step, total, chunk)
    if providers is None or total == 0:
        return None
    stripes = providers[mine["host"]]

    # ---- 3. leaves, plan, staging
    public = _strip_private(meta)
    alloc = _plan.arena_allocator(device, entry.get("host_leaves") or ())

    error = None
    staging = None
    leaves = engine = None
    t_staged = t_begin
    try:
        leaves = _plan.pair_or_alloc(public, into, alloc)
        engine = handler._engine_for(leaves)
        if engine is None:
            raise ValueError("peer restore needs CUDA tensors to restore into")
        if engine.chunk_bytes != chunk:
            raise ValueError("image was checksummed in %d-byte chunks, engine uses %d"
                             % (chunk, engine.chunk_bytes))
        staging = engine.prepare_restore(leaves, total)
        # ---- 4. make my own stripe complete in my staging
        for lo, hi, provider, needs_staging in stripes:
            if provider == me and needs_staging and hi > lo:
                engine.stage_range(handler.shared_memory.buf, total, lo, hi)
                handler._staged = (step, lo, hi)
                handler._buffer_size = total
        t_staged = time.perf_counter()
    except Exception as e:  # noqa: BLE001 - reported collectively below
        error = e

    # ---- 5. exchange staging handles, open the peers' (same host only)
    serving = error is None and any(p == me and hi > lo for lo, hi, p, _ in stripes)
    exported = None
    if serving:
        torch.cuda.synchronize(dev)
        exported = _native.ipc_export(staging.data_ptr())
    handles = [None] * world
    dist.all_gather_object(handles, exported, group=group)
    bases = []
    try:
        if error is None:
            for lo, hi, provider, _ in stripes:
                if provider == me or hi <= lo:
                    bases.append(staging.data_ptr())
                    continue
                if handles[provider] is None:
                    raise RuntimeError("rank %d cannot serve its stripe" % provider)
                handle, offset = handles[provider]
                # (mappings of peers' staging are kept by the engine: see peer_base)
                bases.append(engine.peer_base(provider, handle) + offset)
    except Exception as e:  # noqa: BLE001
        error = error or e
    # everybody's stripe is complete and exported before anybody reads
    _vote(error is None, dev, group)
    t_open = time.perf_counter()
    try:
        if error is None:
            n_chunks = -(-total // chunk)
            per_bytes = -(-n_chunks // len(stripes)) * chunk  # as plan_providers cuts them
            expected = checksums_from_meta(entry, total, chunk)
            local = all(p == me for lo, hi, p, _ in stripes if hi > lo)
            # start one stripe past my own: at any moment the ranks of the
            # node then read different peers
            ranks_here = sorted(r for r, f in enumerate(facts) if f["host"] == mine["host"])
            engine.restore_from_slices(total, bases, per_bytes, expected, local=local,
                                       first_slice=ranks_here.index(me) + 1)
            handler.last_stats = dict(engine.stats)
    except Exception as e:  # noqa: BLE001
        error = error or e
    t_done = time.perf_counter()
    # nobody lets go of its staging before everybody is done reading
    all_ok = _vote(error is None, dev, group)
    if error is not None or not all_ok:
        # whatever went wrong, do not offer this staging again: the next
        # attempt starts from the shm image
        handler._staged = (0, 0, 0)
    if error is not None:
        raise error
    if not all_ok:
        raise _native.ChecksumMismatch(-74, "a peer failed to restore step %d from memory" % step)
    pulled = sum(hi - lo for lo, hi, p, _ in stripes if p != me)
    staged = sum(hi - lo for lo, hi, p, s in stripes if p == me and s)
    handler.last_stats.update({
        "restore_path": handler.last_stats.get("restore_path", "slices_nvlink"),
        "peer_bytes_over_nvlink": pulled, "bytes_over_host_link": staged,
        "peer_stage_s": t_staged - t_begin, "peer_exchange_s": t_open - t_staged,
        "peer_scatter_s": t_done - t_open, "peer_total_s": time.perf_counter() - t_begin})
    logger.info("rank %d restored step %d: %.2f GB over NVLink from peers' HBM, %.2f GB over its "
                "host link, scatter %.1f ms", me, step, pulled / 1e9, staged / 1e9,
                (t_done - t_open) * 1e3)
    it = iter(leaves)

    def fill(m):
        return next(it).tensor if isinstance(m, _plan.TensorMeta) else m

    return step, _plan.traverse_state_dict(public, fill)


def _vote(ok, dev, group):
    """all-reduce of a success flag (doubles as the barrier between the
    phases); device-synchronised on both sides."""
    torch.cuda.synchronize(dev)
    flag = torch.tensor([0 if ok else 1], dtype=torch.int32, device=dev)
    dist.all_reduce(flag, group=group)
    return int(flag.item()) == 0


def applicable(engine, into, device):
    """Replicated state, NCCL job, CUDA targets, not switched off."""
    if os.getenv("FCP_PEER_RESTORE", "1") != "1":
        return False
    if not dist.is_initialized() or dist.get_world_size() < 2:
        return False
    if engine.get_global_shard_num() != 1:
        return False
    if not torch.cuda.is_available() or dist.get_backend() != "nccl":
        return False
    if device is not None:
        return torch.device(device).type == "cuda"
    if into is None:
        return False
    found = []

    def visit(v):
        if torch.is_tensor(v) and v.is_cuda:
            found.append(1)
        return v

    _plan.traverse_state_dict(into, visit)
    return bool(found)
