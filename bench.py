"""Flash-checkpoint bench: one "step" = one save of a rank's shard from HBM into
host shared memory (durable, CRC-32C per 2 MiB chunk) + one restore of it back
into the live tensors (CRC verified).  See DESIGN.md §7 for the definitions.

    python bench.py --gpus 1 --steps 5 --warmup 3
    torchrun --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference's own CPU path

Besides the timed legs (`value`, `e2e`) the default run exercises, outside the
timed regions and with parity asserted in each (a failure exits non-zero):
  N = 1 : `megatron`   MegatronCheckpointer on the config-4 shard (GPT 13B TP=8)
  N >= 2: `fsdp`       FsdpCheckpointer on real FSDP(GPT-2 xl) (config 2)
          `replicated` DdpCheckpointer on replicated state: striped save, restore
                       with stripes over NVLink vs the reference policy
          `recovery`   a rank that lost its state restored from its peers' HBM
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import mmap
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

GB = 1e9
PCIE_GEN5_X16_GBS = 64.0
# dram__bytes_read.sum + dram__bytes_write.sum over the pack launches of one save
# (ncu --set full capture of this program, profiles/r01_ncu_pack_v3_summary.md)
NCU_TRAFFIC_BYTES = {("gpt2-xl-1.5B-adamw-fp32", 0): 39.10e9,      # profiles/r01_ncu_pack_v3_summary.md
                     ("llama2-7b-fsdp8-shard", 0): 24.204e9}       # profiles/r02_ncu_pack_llama_summary.md
NCU_TRAFFIC_SOURCE = "profiles/r02_ncu_pack_llama_summary.md, r01_ncu_pack_v3_summary.md"
_REAL_STDOUT = sys.stdout

WORKLOADS = {
    # nanoGPT shapes, bias=False, block_size 128: fp32 params (tied
    # wte/lm_head stored twice, as the reference does) + AdamW step/exp_avg/exp_avg_sq
    "gpt2-xl-1.5B-adamw-fp32": dict(n_layer=48, n_embd=1600, vocab=50304, block=128),
    "gpt2-small-124M-adamw-fp32": dict(n_layer=12, n_embd=768, vocab=50304, block=128),
}
# sharded workloads (configs 3 and 4 of BASELINE.json): per-rank shard fixed at
# the 8-way size whatever --gpus is
SHARDED = {
    # Llama-2 7B, FSDP over 8 ranks: bf16 param shard + fp32 master / exp_avg /
    # exp_avg_sq + step per tensor  (11.8 GB per rank)
    "llama2-7b-fsdp8-shard": dict(kind="llama", layers=32, hidden=4096, inter=11008,
                                  vocab=32000, ways=8),
    # Megatron GPT 13B, TP=8: fp16 model shard + fp32 master / exp_avg /
    # exp_avg_sq, CPU rng states  (~22.5 GB per rank)
    "gpt-13b-tp8-shard": dict(kind="megatron", layers=40, hidden=5120, vocab=50304, ways=8),
}


def sharded_shapes(workload):
    c = SHARDED[workload]
    h, w = c["hidden"], c["ways"]
    out = []
    if c["kind"] == "llama":
        out.append(("model.embed_tokens.weight", c["vocab"] * h))
        for i in range(c["layers"]):
            p = "model.layers.%d." % i
            for n in ("self_attn.q_proj", "self_attn.k_proj", "self_attn.v_proj",
                      "self_attn.o_proj"):
                out.append((p + n + ".weight", h * h))
            for n in ("mlp.gate_proj", "mlp.up_proj", "mlp.down_proj"):
                out.append((p + n + ".weight", h * c["inter"]))
            out.append((p + "input_layernorm.weight", h))
            out.append((p + "post_attention_layernorm.weight", h))
        out.append(("model.norm.weight", h))
        out.append(("lm_head.weight", c["vocab"] * h))
        return [(n, (numel + w - 1) // w) for n, numel in out]  # FSDP: 1/ways of each
    # Megatron TP: matrices split w ways, layernorms / biases of row-parallel layers whole
    out.append(("embedding.word_embeddings.weight", c["vocab"] * h // w))
    out.append(("embedding.position_embeddings.weight", 2048 * h))
    for i in range(c["layers"]):
        p = "encoder.layers.%d." % i
        out += [(p + "input_layernorm.weight", h), (p + "input_layernorm.bias", h),
                (p + "self_attention.query_key_value.weight", 3 * h * h // w),
                (p + "self_attention.query_key_value.bias", 3 * h // w),
                (p + "self_attention.dense.weight", h * h // w),
                (p + "self_attention.dense.bias", h),
                (p + "post_attention_layernorm.weight", h),
                (p + "post_attention_layernorm.bias", h),
                (p + "mlp.dense_h_to_4h.weight", 4 * h * h // w),
                (p + "mlp.dense_h_to_4h.bias", 4 * h // w),
                (p + "mlp.dense_4h_to_h.weight", 4 * h * h // w),
                (p + "mlp.dense_4h_to_h.bias", h)]
    out += [("encoder.final_layernorm.weight", h), ("encoder.final_layernorm.bias", h)]
    return out


def build_sharded_state(workload, device, seed, max_bytes=None):
    """Per-rank shard of configs 3 / 4 as the checkpointers hand it to the
    engine: low-precision model shard + fp32 optimizer state (+ CPU rng
    states for the Megatron one).  `max_bytes` truncates the tensor list (the
    bounded CPU-baseline sample)."""
    c = SHARDED[workload]
    lowp = torch.bfloat16 if c["kind"] == "llama" else torch.float16
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    model, state, used = {}, {}, 0
    for idx, (name, numel) in enumerate(sharded_shapes(workload)):
        if max_bytes is not None and model and used + numel * 14 + 4 > max_bytes:
            continue  # skip what does not fit, keep the small tensors that do
        used += numel * 14 + 4
        master = torch.empty(numel, dtype=torch.float32, device=device).normal_(0, 0.02, generator=g)
        model[name] = master.to(lowp)
        state[idx] = {"step": torch.tensor(float(idx), device=device), "master": master,
                      "exp_avg": torch.empty_like(master).normal_(0, 1e-3, generator=g),
                      "exp_avg_sq": torch.empty_like(master).uniform_(0, 1e-6, generator=g)}
    sd = {"model": model, "optimizer": {"state": state}, "iteration": 1000}
    if c["kind"] == "megatron":
        sd["rng_state"] = [{"torch_rng_state": torch.get_rng_state(),
                            "cuda_rng_state": torch.cuda.get_rng_state(device)
                            if torch.device(device).type == "cuda" else torch.get_rng_state()}]
    return sd


def gpt2_param_shapes(n_layer, n_embd, vocab, block):
    e = n_embd
    shapes = [("transformer.wte.weight", (vocab, e)), ("transformer.wpe.weight", (block, e))]
    for i in range(n_layer):
        p = "transformer.h.%d." % i
        shapes += [(p + "ln_1.weight", (e,)), (p + "attn.c_attn.weight", (3 * e, e)),
                   (p + "attn.c_proj.weight", (e, e)), (p + "ln_2.weight", (e,)),
                   (p + "mlp.c_fc.weight", (4 * e, e)), (p + "mlp.c_proj.weight", (e, 4 * e))]
    shapes += [("transformer.ln_f.weight", (e,)), ("lm_head.weight", (vocab, e))]
    return shapes


def build_state(workload, device, seed, max_bytes=None):
    """{"model", "optimizer", "step"} as examples/pytorch/nanogpt/train.py:377-394
    builds it; synthetic random-init values.  `max_bytes` truncates the layer
    list (used for the bounded CPU-baseline sample)."""
    if workload in SHARDED:
        return build_sharded_state(workload, device, seed, max_bytes)
    cfg = WORKLOADS[workload]
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    shapes = gpt2_param_shapes(**cfg)
    model, state, used = {}, {}, 0
    tied = None
    idx = 0
    for name, shape in shapes:
        numel = int(np.prod(shape))
        if max_bytes is not None and used + numel * 12 > max_bytes and model:
            break
        if name == "lm_head.weight" and tied is not None:
            model[name] = tied  # tied weight: same tensor, laid out twice
            used += numel * 4
            continue
        w = torch.empty(shape, dtype=torch.float32, device=device).normal_(0, 0.02, generator=g)
        model[name] = w
        if name == "transformer.wte.weight":
            tied = w
        state[idx] = {
            "step": torch.tensor(float(1000 + idx), device=device),
            "exp_avg": torch.empty_like(w).normal_(0, 1e-3, generator=g),
            "exp_avg_sq": torch.empty_like(w).uniform_(0, 1e-6, generator=g),
        }
        idx += 1
        used += numel * 12 + 4
    optim = {"state": state, "param_groups": [
        {"lr": 6e-4, "betas": (0.9, 0.95), "eps": 1e-8, "weight_decay": 0.1,
         "amsgrad": False, "fused": True, "params": list(range(idx))}]}
    return {"model": model, "optimizer": optim, "step": 1000}


def workload_bytes(workload):
    """Payload bytes of the full state, from the shapes alone."""
    if workload in SHARDED:
        lowp = 2
        tensors = sum(n * (lowp + 12) + 4 for _, n in sharded_shapes(workload))
        return tensors + (5056 + 16 if SHARDED[workload]["kind"] == "megatron" else 0)
    total = 0
    for name, shape in gpt2_param_shapes(**WORKLOADS[workload]):
        n = int(np.prod(shape))
        total += n * 4 if name == "lm_head.weight" else n * 12 + 4
    return total


def state_bytes(sd):
    seen = 0

    def walk(v):
        nonlocal seen
        if isinstance(v, dict):
            for x in v.values():
                walk(x)
        elif isinstance(v, list):
            for x in v:
                walk(x)
        elif torch.is_tensor(v):
            seen += v.numel() * v.element_size()

    walk(sd)
    return seen


class ClockSampler(object):
    """nvidia-smi clocks + throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:  # noqa: BLE001
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 6:
                    continue
                try:
                    sm.append(float(f[0]))
                    mx.append(float(f[1]))
                except ValueError:
                    continue
                for n, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            os.unlink(self.path)
        except Exception:  # noqa: BLE001
            pass
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx),
                       reasons=sorted(reasons), samples=len(sm))
        return out


def host_info():
    info = {"cpus": os.cpu_count()}
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemTotal"):
                info["mem_total_gb"] = round(int(line.split()[1]) / 1e6, 1)
            if line.startswith("MemAvailable"):
                info["mem_available_gb"] = round(int(line.split()[1]) / 1e6, 1)
        nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")]
        info["numa_nodes"] = len(nodes)
    except Exception:  # noqa: BLE001
        pass
    return info


# ------------------------------------------------- the reference's own path
def load_reference(run_id=None):
    """-> (reference ckpt_saver module or None, kind, why-not).  The reference
    is Python: its import closure is shipped unmodified in baseline/_ref/ by
    __graft_entry__.build(); the oracle port is the fallback."""
    try:
        from oracle import reference_closure
        return reference_closure.import_reference(run_id), "reference", None
    except Exception as e:  # noqa: BLE001
        return None, "port", "%s: %s" % (type(e).__name__, e)


def _copy_back(views, live):
    """What module / optimizer.load_state_dict do with the reference's views:
    one copy_ per tensor into the live tensor."""
    if isinstance(views, dict):
        for k, v in views.items():
            if isinstance(v, (dict, list)) or torch.is_tensor(v):
                _copy_back(v, live[k])
    elif isinstance(views, list):
        for i, v in enumerate(views):
            if isinstance(v, (dict, list)) or torch.is_tensor(v):
                _copy_back(v, live[i])
    elif views.numel():
        with torch.no_grad():
            live.copy_(views)


class ReferenceLoop(object):
    """One rank's reference path on `sd` (host or CUDA tensors):
    ``SharedMemoryHandler.save_state_dict`` (ckpt_saver.py:272-290: meta walk,
    two SharedDict.set round trips, one ``copy_`` per tensor into the shm) and
    ``load_state_dict`` (:292-311, zero-copy views) followed by one ``copy_``
    per tensor back into the live tensors.  `ref` None: the oracle port of the
    two loops (no meta IPC)."""

    def __init__(self, ref, sd, shard_id):
        self.ref, self.sd = ref, sd
        if ref is not None:
            self.handler = ref.SharedMemoryHandler(shard_id, host=True)
            self.key = ref.DLROVER_CKPT_CONFIG_KEY
        else:
            from oracle import oracle
            self.oracle = oracle
            self.meta, total = oracle.layout(sd)
            self.mm = mmap.mmap(-1, max(total, 1), flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)

    def step(self, k):
        if self.ref is None:
            self.oracle.ref_save_into(self.sd, self.meta, self.mm)
            self.oracle.ref_load_from(self.meta, self.mm, self.sd)
            return
        wrapped = {"model_states": self.sd,
                   self.key: self.ref.CheckpointConfig(step=k, paths={"model_states": "x"})}
        self.handler.save_state_dict(wrapped)
        views = self.handler.load_state_dict()
        _copy_back(views["model_states"], self.sd)

    def close(self):
        if self.ref is None:
            self.mm.close()
        else:
            self.handler.unlink()
            self.handler.close()


def reference_loop_on_gpu_tensors(sd, ref, kind, shard_id, device):
    """The true "before" on GPU-resident state: the reference's handler fed
    the CUDA state dict itself — one synchronous pageable D2H per tensor on
    save, one pageable H2D per tensor on restore.  All ranks at once, the
    whole shard, 1 warm-up + 2 timed steps, max over ranks."""
    loop = ReferenceLoop(ref, sd, shard_id)
    B = state_bytes(sd)
    loop.step(1)
    times = []
    for k in range(2):
        _barrier_sync(device)
        t0 = time.perf_counter()
        loop.step(2 + k)
        torch.cuda.synchronize(device)
        times.append(_dist_max(time.perf_counter() - t0, device))
    loop.close()
    dt = min(times)
    world = dist.get_world_size() if dist.is_initialized() else 1
    return {"value": world * 2 * B / dt / GB, "unit": "GB/s", "s_per_step": dt, "kind": kind,
            "per_rank_GBps": 2 * B / dt / GB, "sample": "the bench shard itself, all ranks at once",
            "what": "reference SharedMemoryHandler.save_state_dict / load_state_dict + per-tensor "
                    "copy_ back, CUDA tensors <-> pageable shm (ckpt_saver.py:272-311)"}


def cpu_baseline_sample(workload, ref, kind, sample_bytes, shard_id):
    """cpu_baseline: the reference path on a bounded host-resident prefix of
    the workload, rank 0 at N=1, with the threads torch picks."""
    sd = build_state(workload, "cpu", 1234, max_bytes=sample_bytes)
    loop = ReferenceLoop(ref, sd, shard_id)
    B = state_bytes(sd)
    loop.step(1)
    t0 = time.perf_counter()
    for k in range(3):
        loop.step(2 + k)
    dt = (time.perf_counter() - t0) / 3
    loop.close()
    n = sum(1 for _ in _iter_tensors(sd))
    return {"value": 2 * B / dt / GB, "unit": "GB/s", "cores": torch.get_num_threads(),
            "kind": kind, "s_per_step": dt,
            "sample": "%s: %d tensors, %.2f GB of host tensors, save+restore" % (workload, n, B / GB)}


# ------------------------------------------------------------------ helpers
def _dist_max(x, device):
    if not dist.is_initialized():
        return x
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _barrier_sync(device):
    if dist.is_initialized():
        dist.barrier()
    if torch.device(device).type == "cuda":
        torch.cuda.synchronize(device)


def measure_host_link(device, nbytes=1 << 30):
    """Copy-engine rate of this rank's link with every rank copying at the same
    time: one pinned (cudaHostAlloc) buffer, device->host and host->device,
    CUDA events, max over ranks.  No kernels of this repo: it is the ceiling
    the save / restore rates are read against."""
    dev = torch.empty(nbytes, dtype=torch.uint8, device=device)
    host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    stream = torch.cuda.current_stream(device)
    out = {}
    for name, (dst, src) in (("d2h", (host, dev)), ("h2d", (dev, host))):
        best = None
        for i in range(4):
            _barrier_sync(device)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            dst.copy_(src, non_blocking=True)
            b.record(stream)
            b.synchronize()
            ms = _dist_max(a.elapsed_time(b), device)
            if i and (best is None or ms < best):
                best = ms
        out[name + "_GBps_per_rank"] = nbytes / (best / 1e3) / GB
    out["how"] = "1 GiB pinned buffer, copy engine, all ranks at once, best of 3, max over ranks"
    del dev, host
    return out


def measure_nvlink(device, rank, world, nbytes=1 << 30):
    """NCCL broadcast of 1 GiB from rank 0 (device-timed, max over ranks): the
    rate peer recovery is read against."""
    buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
    best = None
    for i in range(4):
        _barrier_sync(device)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        dist.broadcast(buf, src=0)
        b.record()
        b.synchronize()
        ms = _dist_max(a.elapsed_time(b), device)
        if i and (best is None or ms < best):
            best = ms
    del buf
    return {"nccl_broadcast_GBps": nbytes / (best / 1e3) / GB,
            "how": "1 GiB dist.broadcast from rank 0, CUDA events, best of 3, max over ranks"}


def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:  # noqa: BLE001
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def _iter_tensors(v):
    if isinstance(v, dict):
        for x in v.values():
            yield from _iter_tensors(x)
    elif isinstance(v, list):
        for x in v:
            yield from _iter_tensors(x)
    elif torch.is_tensor(v):
        yield v


def _wreck(sd):
    with torch.no_grad():
        for t in _iter_tensors(sd):
            if t.is_cuda and t.numel():
                t.reshape(-1).view(torch.uint8)[::4097] ^= 0x5A  # cheap, touches every tensor


def _digest(sd, device):
    """One number per state dict: sum over tensors of the wrapped int64 sum of
    their bytes — cheap to compute on the device and order-sensitive enough to
    notice any tensor that was not restored."""
    acc = torch.zeros((), dtype=torch.int64, device=device)
    for i, t in enumerate(_iter_tensors(sd)):
        if t.is_cuda and t.numel():
            raw = t.reshape(-1).view(torch.uint8)
            acc += raw.sum(dtype=torch.int64) * (i + 1)
    return int(acc.item())


def _check_image_against_oracle(handler, what):
    """Every chunk of the shm image: oracle CRC-32C (crc32 instruction, checked
    against the bytewise statement in tests/test_oracle.py) == recorded sum."""
    from dlrover_wuqiong_b200.ckpt_saver import FCP_CHECKSUM_KEY
    from oracle import oracle
    entry = handler.metadata.get()[FCP_CHECKSUM_KEY]
    got = oracle.chunk_crc32c_fast(handler.shared_memory.buf, entry["chunk_bytes"])
    want = np.asarray(entry["sums"], dtype=np.uint32)
    bad = np.nonzero(got != want)[0]
    assert len(got) == len(want) and bad.size == 0, \
        "%s: chunk %s of the shm image differs from the oracle's CRC-32C" % (what, bad[:4])
    return int(len(got))


# ----------------------------------------------------------- reference arm
def run_reference_arm(args, rank, local_rank, world):
    """--impl reference: the reference's own CPU implementation of the path on
    the host cores — every rank its own full shard at the same time (threads =
    cores / N), exactly what N trainers of the reference would do."""
    if world > 1:
        dist.init_process_group("gloo")
    cores = os.cpu_count() or 1
    torch.set_num_threads(max(1, cores // world))
    ref, kind, why = load_reference("fcpbench_ref_%s" % os.environ.get("MASTER_PORT", os.getpid()))
    workload = args.workload
    hi = host_info()
    B_full = workload_bytes(workload)
    avail = hi.get("mem_available_gb", 1e9) * GB
    sample_bytes = None
    if world * B_full * 2.3 > 0.8 * avail:  # state + shm per rank must fit the host
        sample_bytes = int(0.8 * avail / world / 2.3)
    sd = build_state(workload, "cpu", 1234 + rank, max_bytes=sample_bytes)
    B = state_bytes(sd)
    n_tensors = sum(1 for _ in _iter_tensors(sd))
    first = sd["model"][next(iter(sd["model"]))]
    want = first.clone()
    loop = ReferenceLoop(ref, sd, local_rank)
    for w in range(max(args.warmup, 1)):
        loop.step(1 + w)
    _barrier_sync("cpu")
    t0 = time.perf_counter()
    for k in range(args.steps):
        loop.step(100 + k)
    dt = time.perf_counter() - t0
    _barrier_sync("cpu")
    dt = _dist_max(dt, "cpu") / args.steps
    assert torch.equal(first, want), "reference round trip failed"
    loop.close()
    if rank == 0:
        gbs = world * 2 * B / dt / GB
        sample = "%s: every rank's whole shard%s, %d tensors, %.2f GB per rank, save+restore" % (
            workload, "" if sample_bytes is None else " (truncated to fit host memory)",
            n_tensors, B / GB)
        line = {
            "impl": "reference", "metric": "flash_ckpt_save_plus_restore_GBps", "value": gbs,
            "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload, "shard_bytes_per_rank": B,
                       "tensors_per_rank": n_tensors, "ranks": world,
                       "threads_per_rank": torch.get_num_threads(), "sample": sample,
                       "same_config": sample_bytes is None,
                       "what": "reference SharedMemoryHandler.save_state_dict + load_state_dict "
                               "+ per-tensor copy_ back, host tensors (ckpt_saver.py:272-311)"
                               if ref is not None else
                               "oracle port of ckpt_saver.py:174-206,112-137 (%s)" % why},
            "cpu_baseline": {"value": gbs, "unit": "GB/s",
                             "cores": torch.get_num_threads() * world, "kind": kind,
                             "sample": sample},
            "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "host": hi,
        }
        print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    if dist.is_initialized():
        dist.destroy_process_group()


# -------------------------------------------------------------------- legs
def leg_megatron(device, ckpt_root):
    """Config 4 through its own checkpointer: MegatronCheckpointer on one
    rank's GPT-13B TP=8 shard (22.6 GB, 2422 tensors, host rng states), with
    megatron_save / megatron_load patched the way the reference's own test
    does (megatron_ckpt_test.py:84-115)."""
    import argparse as ap
    from dlrover_wuqiong_b200.flash_checkpoint import megatron as mg
    from dlrover_wuqiong_b200.flash_checkpoint.checkpointer import StorageType
    workload = "gpt-13b-tp8-shard"
    free_hbm, _ = torch.cuda.mem_get_info(device)
    B_est = workload_bytes(workload)
    if 3.3 * B_est > free_hbm:
        return {"skipped": "needs %.0f GB of HBM, %.0f free" % (3.3 * B_est / GB, free_hbm / GB)}
    sd = build_state(workload, device, 4321)
    B = state_bytes(sd)
    save_dir = os.path.join(ckpt_root, "megatron")
    suffix = "model_optim_rng.pt"

    def get_args():
        a = ap.Namespace()
        a.save = save_dir
        return a

    def fake_save(iteration, model, optimizer, opt_param_scheduler, *rest):
        torch.save({"iteration": iteration, "model": model["model"],
                    "optimizer": optimizer["optimizer"], "rng_state": model["rng_state"]},
                   os.path.join(save_dir, "iter_%07d" % iteration, "mp_rank_00", suffix))

    seen = {}

    def fake_load(model, optimizer, opt_param_scheduler, load_arg="load", strict=True):
        got = torch.load(os.path.join(save_dir, "iter_%07d" % seen["iteration"], "mp_rank_00",
                                      suffix))
        seen["devices"] = sorted({t.device.type for t in _iter_tensors(got["model"])})
        seen["rng_device"] = got["rng_state"][0]["torch_rng_state"].device.type
        _copy_back(got["model"], model["model"])            # model.load_state_dict
        _copy_back(got["optimizer"], optimizer["optimizer"])  # optimizer.load_state_dict
        torch.set_rng_state(got["rng_state"][0]["torch_rng_state"])
        return got["iteration"]

    prev = mg.megatron_save, mg.megatron_load, mg.get_args
    mg.megatron_save, mg.megatron_load, mg.get_args = fake_save, fake_load, get_args
    mg.MegatronCheckpointer._instance = None
    out = {"workload": workload, "bytes": B, "tensors": sum(1 for _ in _iter_tensors(sd)),
           "api": "flash_checkpoint.megatron.save_checkpoint(MEMORY) / load_checkpoint"}
    try:
        holder = {"model": sd["model"], "rng_state": sd["rng_state"]}
        opt = {"optimizer": sd["optimizer"]}
        torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        assert mg.save_checkpoint(10, holder, opt, None, storage_type=StorageType.MEMORY)
        out["first_save_s"] = time.perf_counter() - t0
        times = []
        for it in (11, 12):
            torch.cuda.synchronize(device)
            t0 = time.perf_counter()
            assert mg.save_checkpoint(it, holder, opt, None, storage_type=StorageType.MEMORY)
            times.append(time.perf_counter() - t0)
        out["save_s"] = min(times)
        out["save_GBps"] = B / out["save_s"] / GB
        handler = mg.MegatronCheckpointer.singleton_instance(save_dir).engine._shm_handler
        out["chunks_checked_vs_oracle"] = _check_image_against_oracle(handler, "megatron leg")
        want = _digest(sd, device)
        seen["iteration"] = 12
        for name, env in (("load_s", None), ("load_reference_policy_s", "cpu")):
            _wreck(sd)
            assert _digest(sd, device) != want
            if env:
                os.environ["FCP_RESTORE_DEVICE"] = env
            torch.cuda.synchronize(device)
            t0 = time.perf_counter()
            it = mg.load_checkpoint(holder, opt, None)
            torch.cuda.synchronize(device)
            out[name] = time.perf_counter() - t0
            os.environ.pop("FCP_RESTORE_DEVICE", None)
            assert it == 12 and _digest(sd, device) == want, "megatron restore differs"
            if env is None:
                out["load_breakdown"] = {k: handler.last_stats.get(k) for k in (
                    "restore_prepare_s", "restore_move_s", "restore_host_leaves_s")}
                out["restore_path"] = handler.last_stats.get("restore_path")
                out["framework_got_tensors_on"] = seen["devices"]
                assert out["restore_path"] in ("registered", "bounced"), out["restore_path"]
                assert seen["devices"] == ["cuda"] and seen["rng_device"] == "cpu", seen
        out["load_GBps"] = B / out["load_s"] / GB
        out["load_reference_policy"] = ("FCP_RESTORE_DEVICE=cpu: CPU views of the shm + one "
                                        "pageable H2D per tensor, as the reference does")
        out["parity"] = "all chunk CRCs vs oracle; digest of all tensors after both restores"
    finally:
        mg.megatron_save, mg.megatron_load, mg.get_args = prev
        inst = mg.MegatronCheckpointer._instance
        if inst is not None:
            inst.engine.close()
        mg.MegatronCheckpointer._instance = None
    return out


def _gpt2_module(n_layer, n_embd, n_head, vocab, block):
    """nanoGPT's module tree (examples/pytorch/nanogpt/model.py), bias=False,
    tied wte / lm_head — same parameter names and shapes."""
    import torch.nn as nn
    import torch.nn.functional as F

    class Attn(nn.Module):
        def __init__(self):
            super().__init__()
            self.c_attn = nn.Linear(n_embd, 3 * n_embd, bias=False)
            self.c_proj = nn.Linear(n_embd, n_embd, bias=False)

        def forward(self, x):
            b, t, c = x.shape
            q, k, v = self.c_attn(x).split(n_embd, dim=2)
            q, k, v = (z.view(b, t, n_head, c // n_head).transpose(1, 2) for z in (q, k, v))
            y = F.scaled_dot_product_attention(q, k, v, is_causal=True)
            return self.c_proj(y.transpose(1, 2).reshape(b, t, c))

    class MLP(nn.Module):
        def __init__(self):
            super().__init__()
            self.c_fc = nn.Linear(n_embd, 4 * n_embd, bias=False)
            self.c_proj = nn.Linear(4 * n_embd, n_embd, bias=False)

        def forward(self, x):
            return self.c_proj(F.gelu(self.c_fc(x)))

    class Block(nn.Module):
        def __init__(self):
            super().__init__()
            self.ln_1 = nn.LayerNorm(n_embd, bias=False)
            self.attn = Attn()
            self.ln_2 = nn.LayerNorm(n_embd, bias=False)
            self.mlp = MLP()

        def forward(self, x):
            x = x + self.attn(self.ln_1(x))
            return x + self.mlp(self.ln_2(x))

    class GPT(nn.Module):
        def __init__(self):
            super().__init__()
            self.transformer = nn.ModuleDict(dict(
                wte=nn.Embedding(vocab, n_embd), wpe=nn.Embedding(block, n_embd),
                h=nn.ModuleList([Block() for _ in range(n_layer)]),
                ln_f=nn.LayerNorm(n_embd, bias=False)))
            self.lm_head = nn.Linear(n_embd, vocab, bias=False)
            self.transformer.wte.weight = self.lm_head.weight

        def forward(self, idx):
            pos = torch.arange(idx.shape[1], device=idx.device)
            x = self.transformer.wte(idx) + self.transformer.wpe(pos)
            for blk in self.transformer.h:
                x = blk(x)
            return self.lm_head(self.transformer.ln_f(x))

    return GPT(), Block


def leg_fsdp(device, rank, world, ckpt_root, saves=4):
    """Config 2 through its own checkpointer: nanoGPT GPT-2 xl (48 x 1600,
    1.5 B parameters, fp32 + AdamW) wrapped in real FSDP over all ranks, one
    optimizer step, FsdpCheckpointer.save_checkpoint(MEMORY) x saves (first:
    planned; later: cached DCP plan), wreck, load_checkpoint, compare."""
    from torch.distributed.fsdp import FullyShardedDataParallel as FSDP
    from torch.distributed.fsdp.wrap import ModuleWrapPolicy
    from dlrover_wuqiong_b200.flash_checkpoint import fsdp_engine
    from dlrover_wuqiong_b200.flash_checkpoint.fsdp import FsdpCheckpointer, StorageType
    torch.manual_seed(77)
    with torch.device(device):  # (not via the meta device: to_empty would untie wte / lm_head)
        model, Block = _gpt2_module(48, 1600, 16, 50304, 128)
    model = FSDP(model, auto_wrap_policy=ModuleWrapPolicy({Block}), device_id=device)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-4, fused=True)

    def train_step():
        opt.zero_grad(set_to_none=True)
        idx = torch.randint(0, 50304, (2, 128), device=device)
        model(idx).float().square().mean().backward()
        opt.step()

    def live_tensors():
        out = [p for p in model.parameters()]
        for st in opt.state.values():  # by key: load_state_dict may reorder a state's entries
            out += [st[k] for k in sorted(st) if torch.is_tensor(st[k])]
        return out

    train_step()
    ckpt = FsdpCheckpointer(os.path.join(ckpt_root, "fsdp"))
    eng = ckpt._engine
    rows = []
    for k in range(saves):
        if k:
            train_step()
        _barrier_sync(device)
        t0 = time.perf_counter()
        assert ckpt.save_checkpoint(k + 1, model, opt, {"step": k + 1},
                                    storage_type=StorageType.MEMORY)
        total = time.perf_counter() - t0
        rows.append({"planned": bool(eng.last_save_planned),
                     "save_checkpoint_s": _dist_max(total, device),
                     "save_to_memory_s": _dist_max(eng.last_save_to_memory_s, device)})
    assert rows[0]["planned"] and not rows[-1]["planned"], rows
    handler = eng._shm_handler
    shard_bytes = handler.shared_memory.size
    chunks = _check_image_against_oracle(handler, "fsdp leg")
    durable_ms = handler.last_stats.get("total_ms")
    snap = [t.detach().clone() for t in live_tensors()]
    with torch.no_grad():
        for t in live_tensors():
            t.zero_()
    # split load_checkpoint into the byte path (SharedMemoryReader.read_data:
    # H2D + scatter kernel) and the framework around it (DCP planning,
    # state-dict assembly, optimizer re-flattening)
    spent = {"read_data": 0.0}
    real = fsdp_engine.SharedMemoryReader.read_data

    def timed_read(self, plan, planner):
        t = time.perf_counter()
        try:
            return real(self, plan, planner)
        finally:
            torch.cuda.synchronize(device)
            spent["read_data"] += time.perf_counter() - t

    fsdp_engine.SharedMemoryReader.read_data = timed_read
    try:
        _barrier_sync(device)
        t0 = time.perf_counter()
        extra = ckpt.load_checkpoint(model, opt)
        torch.cuda.synchronize(device)
        load_s = time.perf_counter() - t0
    finally:
        fsdp_engine.SharedMemoryReader.read_data = real
    assert extra["step"] == saves, extra
    n_params = sum(1 for _ in model.parameters())
    now = live_tensors()
    assert len(now) == len(snap), "fsdp restore: %d tensors before, %d after" % (len(snap), len(now))
    for i, (a, b) in enumerate(zip(snap, now)):
        if not torch.equal(a, b.detach()):
            bad = (a != b.detach()).reshape(-1).nonzero()
            raise AssertionError(
                "fsdp restore differs: tensor %d (%s, shape %s, %s): %d of %d elements, first at %d"
                % (i, "parameter" if i < n_params else "optimizer state", tuple(a.shape), a.dtype,
                   bad.numel(), a.numel(), int(bad[0]) if bad.numel() else -1))
    restore_path = handler.last_stats.get("restore_path")
    assert restore_path in ("registered", "bounced"), restore_path
    n_items = len(handler.metadata.get(local=True)["dcp_metadata"].storage_data)
    out = {"workload": "gpt2-xl-1.5B FSDP(%d) + AdamW fp32" % world,
           "api": "FsdpCheckpointer.save_checkpoint(MEMORY) / load_checkpoint",
           "shard_bytes_rank0": shard_bytes, "plan_items": n_items, "saves": rows,
           "planned_save_to_memory_s": rows[0]["save_to_memory_s"],
           "cached_plan_save_to_memory_s": min(r["save_to_memory_s"] for r in rows[1:]),
           "durable_device_s": (durable_ms or 0) / 1e3,
           "load_checkpoint_s": _dist_max(load_s, device),
           "load_byte_path_s": _dist_max(spent["read_data"], device),
           "load_framework_s": _dist_max(load_s - spent["read_data"], device),
           "restore_path": restore_path, "chunks_checked_vs_oracle": chunks,
           "parity": "all chunk CRCs of rank 0's segment vs oracle; torch.equal on every "
                     "parameter and optimizer tensor after wreck + load"}
    ckpt.close()
    del snap, model, opt
    torch.cuda.empty_cache()
    return out


def leg_replicated(device, rank, world, ckpt_root, link):
    """Replicated (DDP) state — GPT-2 xl fp32 + AdamW, 18.99 GB, identical on
    every rank — through DdpCheckpointer: the reference's policy (local rank 0
    writes the whole image; every rank reads the whole image back through the
    host) against striped saves and restores whose stripes travel over NVLink.
    Also the `recovery` numbers: one rank loses its state and gets it back
    from its peers' HBM through load_checkpoint."""
    from dlrover_wuqiong_b200.ckpt_saver import FCP_CHECKSUM_KEY
    from dlrover_wuqiong_b200.flash_checkpoint.ddp import DdpCheckpointer, StorageType
    workload = "gpt2-xl-1.5B-adamw-fp32"
    free_hbm, _ = torch.cuda.mem_get_info(device)
    if 2.3 * workload_bytes(workload) > free_hbm:
        workload = "gpt2-small-124M-adamw-fp32"
    sd = build_state(workload, device, 999)  # same seed on every rank: replicas
    B = state_bytes(sd)
    want = _digest(sd, device)
    nv = measure_nvlink(device, rank, world)
    rep, rec = {"workload": workload, "bytes": B, "ranks": world}, {}

    def timed(fn):
        _barrier_sync(device)
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize(device)
        return r, _dist_max(time.perf_counter() - t0, device)

    def restored(what):
        ok = _digest(sd, device) == want
        flag = torch.tensor([0 if ok else 1], device=device)
        dist.all_reduce(flag)
        assert int(flag.item()) == 0, "%s: restored state differs" % what

    # ---- (1) the reference's policy: one writer, everybody reads everything
    single = DdpCheckpointer(os.path.join(ckpt_root, "rep_single"))
    h = single._engine._shm_handler
    single.save_checkpoint(1, sd, storage_type=StorageType.MEMORY)
    _, t = timed(lambda: single.save_checkpoint(2, sd, storage_type=StorageType.MEMORY))
    rep["single_writer_save_s"] = t
    sums_single = None
    if rank == 0:
        rep["chunks_checked_vs_oracle"] = _check_image_against_oracle(h, "replicated leg")
        sums_single = list(h.metadata.get()[FCP_CHECKSUM_KEY]["sums"])
    # recovery A: rank world-1 lost its tensors, rank 0 holds the whole image in HBM
    if rank == world - 1:
        _wreck(sd)
    _, t = timed(lambda: single.load_checkpoint(into=sd))
    restored("recovery from one survivor")
    st = h.last_stats
    rec["from_one_survivor"] = {
        "s": t, "GBps_into_the_lost_rank": B / t / GB,
        "bytes_over_nvlink": _dist_max(float(st.get("peer_bytes_over_nvlink", 0)), device),
        "bytes_over_host_link": _dist_max(float(st.get("bytes_over_host_link", 0)), device),
        "scatter_over_nvlink_s": _dist_max(st.get("peer_scatter_s", 0.0), device)}
    # everybody restarted: nothing in HBM.  Reference policy first (peer path off)
    os.environ["FCP_PEER_RESTORE"] = "0"
    h._drop_engine()
    _wreck(sd)
    _, t = timed(lambda: single.load_checkpoint(into=sd))
    restored("replicated restore, every rank reads the shm")
    rep["restore_every_rank_reads_shm_s"] = t
    rep["restore_every_rank_reads_shm_host_bytes"] = world * B
    os.environ["FCP_PEER_RESTORE"] = "1"
    h._drop_engine()
    _wreck(sd)
    _, t = timed(lambda: single.load_checkpoint(into=sd))
    restored("replicated restore, stripes over NVLink")
    rep["restore_striped_s"] = t
    rep["restore_striped_host_bytes"] = B
    rep["restore_striped_breakdown"] = {
        k: _dist_max(h.last_stats.get(k, 0.0), device)
        for k in ("peer_stage_s", "peer_exchange_s", "peer_scatter_s")}
    assert h.last_stats.get("restore_path") in ("slices_nvlink", "slices_local")
    single.close()
    dist.barrier()

    # ---- (2) striped save: every rank writes 1/n of the one image
    striped = DdpCheckpointer(os.path.join(ckpt_root, "rep_striped"), stripe_replicas=True)
    hs = striped._engine._shm_handler
    striped.save_checkpoint(3, sd, storage_type=StorageType.MEMORY)
    _, t = timed(lambda: striped.save_checkpoint(4, sd, storage_type=StorageType.MEMORY))
    rep["striped_save_s"] = t
    rep["striped_save_GBps_into_host"] = B / t / GB
    rep["single_writer_save_GBps_into_host"] = B / rep["single_writer_save_s"] / GB
    dist.barrier()
    if rank == 0:  # bit-identical to the single-writer image: same sums, and they verify
        assert list(hs.metadata.get()[FCP_CHECKSUM_KEY]["sums"]) == sums_single, \
            "striped image differs from the single-writer image"
        _check_image_against_oracle(hs, "striped save")
    # recovery B: after a striped save every rank holds its stripe in HBM; one
    # rank lost its process state (tensors and staging)
    if rank == world - 1:
        hs._drop_engine()
        _wreck(sd)
    _, t = timed(lambda: striped.load_checkpoint(into=sd))
    restored("recovery from the peers' stripes")
    st = hs.last_stats
    rec["from_peers_stripes"] = {
        "s": t, "GBps_into_the_lost_rank": B / t / GB,
        "bytes_over_nvlink": _dist_max(float(st.get("peer_bytes_over_nvlink", 0)), device),
        "bytes_over_host_link": _dist_max(float(st.get("bytes_over_host_link", 0)), device)}
    rec["nvlink_measured"] = nv
    rec["api"] = "DdpCheckpointer.load_checkpoint(into=state_dict); CRC-32C of every chunk verified"
    rep["speedup_restore"] = rep["restore_every_rank_reads_shm_s"] / rep["restore_striped_s"]
    rep["speedup_save"] = rep["single_writer_save_s"] / rep["striped_save_s"]
    rep["host_link_measured"] = link
    rep["parity"] = ("single-writer image: every chunk CRC vs oracle; striped image: same sums; "
                     "digest of all tensors on every rank after each restore")
    striped.close()
    del sd
    torch.cuda.empty_cache()
    return rep, rec


# ------------------------------------------------------------- product arm
def run_ours(args, rank, local_rank, world):
    from dlrover_wuqiong_b200 import _native
    from dlrover_wuqiong_b200.ckpt_saver import (
        DLROVER_CKPT_CONFIG_KEY, AsyncCheckpointSaver, CheckpointConfig, SharedMemoryHandler)
    from dlrover_wuqiong_b200.common.multi_process import clear_sock_dir
    from dlrover_wuqiong_b200.flash_checkpoint.ddp import DdpCheckpointer, StorageType

    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    ref, ref_kind, ref_why = load_reference()

    # ---- size guards (a box driven out of memory is a strike)
    hi = host_info()
    workload = args.workload
    B = workload_bytes(workload)
    avail = hi.get("mem_available_gb", 1e9) * GB
    need_host = world * B * 2.3 + (2.5 * args.cpu_sample_gb * GB if world == 1 else 0)
    free_hbm, _ = torch.cuda.mem_get_info(device)
    note = None
    try:
        st = os.statvfs("/dev/shm")
        shm_free = st.f_bavail * st.f_frsize
    except OSError:
        shm_free = avail
    if need_host > 0.85 * avail or world * B * 2.1 > shm_free or 2.3 * B > free_hbm:
        note = "%s does not fit (host %.0f GB avail, HBM %.0f GB free)" % (
            workload, avail / GB, free_hbm / GB)
        workload = "gpt2-small-124M-adamw-fp32"
        B = workload_bytes(workload)

    link = measure_host_link(device)
    sd = build_state(workload, device, 1234 + rank)
    B = state_bytes(sd)
    n_tensors = sum(1 for _ in _iter_tensors(sd))

    ckpt_dir = os.path.join(tempfile.gettempdir(), "fcpbench_%s" % os.environ["TORCHELASTIC_RUN_ID"])
    if local_rank == 0:
        os.makedirs(ckpt_dir, exist_ok=True)
        AsyncCheckpointSaver.start_async_saving_ckpt()
    ckpt = DdpCheckpointer(ckpt_dir, local_shard_num=world, global_shard_num=world)
    handler = ckpt._engine._shm_handler
    path = os.path.join(ckpt_dir, "bench", "rank_%d.pt" % rank)

    def e2e_step(step):
        ok = ckpt.save_checkpoint(step, sd, path, storage_type=StorageType.MEMORY)
        assert ok, "save skipped"
        ckpt.load_checkpoint(into=sd)

    # ---- warm-up through the public API (first save creates + pins the shm)
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    ckpt.save_checkpoint(1, sd, path, storage_type=StorageType.MEMORY)
    first_save_s = time.perf_counter() - t0
    register_s = handler.last_stats.get("register_s")
    for w in range(max(args.warmup, 3)):
        e2e_step(2 + w)

    # ---- parity at full size: EVERY chunk of the shm image against the
    # oracle's CRC-32C, and a wrecked state restored to the same digest
    parity = {"chunks_checked": 0, "tensors_checked": 0}
    if rank == 0:
        parity["chunks_checked"] = _check_image_against_oracle(handler, "main leg")
    want = _digest(sd, device)
    _wreck(sd)
    assert _digest(sd, device) != want
    ckpt.load_checkpoint(into=sd)
    assert _digest(sd, device) == want, "restore differs from what was saved"
    parity["tensors_checked"] = n_tensors
    parity["how"] = ("every 2 MiB chunk of rank 0's image vs the oracle's CRC-32C; every rank: "
                     "all tensors perturbed, restored, byte digest equal")

    # ---- native handles for the device-timed leg
    eng = handler._engine
    plan, mode = eng._plan, eng.mode
    staging_ptr = eng._staging.data_ptr() if eng._staging is not None else None
    addr = eng._mapping.addr
    stream = torch.cuda.current_stream(device)
    K = args.steps
    launches = 0

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()

    # ---- leg 1: `value` — C ABI, inputs resident in HBM, device-timed
    save_ms, pack_ms, restore_ms = [], [], []
    _barrier_sync(device)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    gate = eng._gate  # nodes with > 4 local ranks: device->host copies take turns (DmaGate)
    plan.set_defer_copies(gate.enabled)
    for _ in range(K):
        plan.save_async(staging_ptr, addr, mode, _native.SUM_CRC32C, stream.cuda_stream)
        if gate.enabled:
            gate.acquire()
            plan.release_copies()
        step_sums = plan.save_wait()
        gate.release()
        launches += plan.last_launches()
        t = plan.timings()
        save_ms.append(t["total_ms"])
        pack_ms.append(t["kernel_ms"])
        plan.restore(addr, staging_ptr, mode, _native.SUM_CRC32C, step_sums, stream.cuda_stream)
        launches += plan.last_launches()
        restore_ms.append(plan.timings()["total_ms"])
    ev1.record(stream)
    _barrier_sync(device)
    dev_ms = _dist_max(ev0.elapsed_time(ev1), device)

    # ---- leg 2: `e2e` — DdpCheckpointer.save_checkpoint(MEMORY) + load_checkpoint(into)
    _barrier_sync(device)
    e2e_prof = None
    if os.getenv("FCP_PROFILE_E2E") and rank == 0:  # tuning aid: where the host time goes
        import cProfile
        e2e_prof = cProfile.Profile()
        e2e_prof.enable()
    t0 = time.perf_counter()
    for k in range(K):
        e2e_step(100 + k)
    torch.cuda.synchronize(device)
    e2e_s = _dist_max(time.perf_counter() - t0, device)
    if e2e_prof:
        import pstats
        e2e_prof.disable()
        pstats.Stats(e2e_prof, stream=sys.stderr).sort_stats("tottime").print_stats(30)
    _barrier_sync(device)

    # ---- leg 3: training-blocking time of a save through the Python API
    block_s = []
    for k in range(K):
        wrapped = {"model_states": sd,
                   DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=200 + k, paths={"model_states": path})}
        torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        pending = handler.save_state_dict_async(wrapped)
        pending.wait_packed()
        block_s.append(time.perf_counter() - t0)
        pending.wait()
    clocks = sampler.stop() if rank == 0 else None

    # ---- leg 4: cold restore — what a restarted trainer sees: attach to the
    # existing segment (not pinned in this "new" process) and restore
    handler._drop_engine()  # forget plan, staging and the pinned mapping
    _barrier_sync(device)
    fresh = SharedMemoryHandler(ckpt._engine.local_shard_id, host=False)
    t0 = time.perf_counter()
    fresh.load_state_dict(into={"model_states": sd})
    torch.cuda.synchronize(device)
    restore_cold_s = _dist_max(time.perf_counter() - t0, device)
    cold_path = fresh.last_stats.get("restore_path")
    cold_device_ms = fresh.last_stats.get("total_ms")
    cold_breakdown = {k: fresh.last_stats.get(k) for k in
                      ("restore_prepare_s", "restore_move_s", "restore_host_leaves_s")}
    cold_breakdown["attach_and_meta_s"] = restore_cold_s - sum(
        v or 0 for v in cold_breakdown.values())
    fresh.close()
    _barrier_sync(device)

    save_blocking_s = _dist_max(statistics.median(block_s), device)
    save_pack_ms = _dist_max(statistics.median(pack_ms), device)
    save_durable_ms = _dist_max(statistics.median(save_ms), device)
    restore_med_ms = _dist_max(statistics.median(restore_ms), device)
    ms_per_step = dev_ms / K
    value = world * 2 * B / (ms_per_step / 1e3) / GB
    e2e_value = world * 2 * B / (e2e_s / K) / GB

    # ---- the reference's own code beside it (outside the timed regions)
    ref_gpu = cpu_base = None
    if not args.no_cpu_baseline:
        ref_gpu = reference_loop_on_gpu_tensors(sd, ref, ref_kind, 100 + local_rank, device)
        if rank == 0 and world == 1:
            cpu_base = cpu_baseline_sample(workload, ref, ref_kind,
                                           int(args.cpu_sample_gb * GB), 200)

    # ---- free the main leg's memory, then the per-config legs
    ckpt.close()
    del sd
    torch.cuda.empty_cache()
    legs = {}
    want_legs = set(args.legs.split(",")) if args.legs != "auto" else (
        {"megatron"} if world == 1 else {"fsdp", "replicated"})
    if args.legs == "none" or note is not None:
        want_legs = set()
    if "megatron" in want_legs and world == 1:
        legs["megatron"] = leg_megatron(device, ckpt_dir)
    if "fsdp" in want_legs and world > 1:
        legs["fsdp"] = leg_fsdp(device, rank, world, ckpt_dir)
    if "replicated" in want_legs and world > 1:
        legs["replicated"], legs["recovery"] = leg_replicated(device, rank, world, ckpt_dir, link)

    if rank == 0:
        hbm_peak, peak_src = _peaks()
        achieved = 2 * B / (statistics.mean(pack_ms) / 1e3) / GB
        line = {
            "metric": "flash_ckpt_save_plus_restore_GBps", "value": value, "unit": "GB/s",
            "n_gpus": world, "steps": K, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {
                "workload": workload, "shard_bytes_per_rank": B, "tensors_per_rank": n_tensors,
                "step": "save HBM->shm (durable, CRC32C/2MiB) + restore shm->live tensors (verified)",
                "transfer_mode": {0: "staged", 1: "direct", 2: "staged_kernel"}[mode],
                "transfer_mode_chosen_at_run_time": (
                    dict(eng.tuner.rates, state=eng.tuner.state) if eng.tuner is not None else None),
                "l2": "inputs (%.1f GB) larger than L2" % (B / GB), "note": note,
                "dma_writers_at_a_time": (int(os.getenv("FCP_DMA_WRITERS", "4"))
                                          if gate.enabled else world),
                "host_link_peak_GBps_per_dir": PCIE_GEN5_X16_GBS},
            "save_blocking_s": save_blocking_s,
            "save_blocking_device_s": save_pack_ms / 1e3,
            "save_durable_s": save_durable_ms / 1e3,
            "restore_s": restore_med_ms / 1e3,
            "restore_cold_s": restore_cold_s, "restore_cold_path": cold_path,
            "restore_cold_copy_and_scatter_s": (cold_device_ms or 0) / 1e3,
            "restore_cold_breakdown": cold_breakdown,
            "restore_cold_GBps_per_rank": B / restore_cold_s / GB,
            "save_GBps_per_rank": B / (save_durable_ms / 1e3) / GB,
            "restore_GBps_per_rank": B / (restore_med_ms / 1e3) / GB,
            "host_link_frac": {"save": B / (save_durable_ms / 1e3) / GB / PCIE_GEN5_X16_GBS,
                               "restore": B / (restore_med_ms / 1e3) / GB / PCIE_GEN5_X16_GBS},
            "host_link_measured": link,
            "host_link_frac_of_measured": {
                "save": B / (save_durable_ms / 1e3) / GB / link["d2h_GBps_per_rank"],
                "restore": B / (restore_med_ms / 1e3) / GB / link["h2d_GBps_per_rank"]},
            "first_save_s": first_save_s, "host_register_s": register_s,
            "roofline": {"bound": "hbm", "kernel": "fcp_kernel<PACK,CRC,32,1024> (pack + fused CRC32C)",
                         "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                         "frac": achieved / hbm_peak,
                         "traffic": NCU_TRAFFIC_BYTES.get((workload, mode)),
                         "traffic_unit": "bytes per launch set (dram read + write of one save's "
                                         "pack launches, ncu --set full, %s)" % NCU_TRAFFIC_SOURCE,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch_set": 2 * B},
            "cpu_baseline": cpu_base,
            "reference_loop_gpu_source": ref_gpu,
            "reference_import": {"kind": ref_kind, "why_not": ref_why},
            "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": B,
                    "d2h_bytes_per_step": B, "s_per_step": e2e_s / K,
                    "api": "DdpCheckpointer.save_checkpoint(MEMORY) + load_checkpoint(into=)"},
            "gpu_launches": launches,
            "clocks": clocks, "parity": parity, "host": hi,
        }
        line.update(legs)
        print(json.dumps(line), file=_REAL_STDOUT, flush=True)

    if dist.is_initialized():
        dist.barrier()
    if local_rank == 0:
        saver = AsyncCheckpointSaver.get_ckpt_saver()
        if saver is not None:
            saver.close()
        clear_sock_dir()
    if dist.is_initialized():
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="llama2-7b-fsdp8-shard",
                    choices=sorted(WORKLOADS) + sorted(SHARDED))
    ap.add_argument("--cpu-sample-gb", type=float, default=4.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--legs", default="auto",
                    help="auto | none | comma list of megatron,fsdp,replicated")
    args = ap.parse_args()

    # keep stdout for the one JSON line: libraries (NCCL's version banner, ...)
    # write to fd 1 behind Python's back, so point fd 1 at stderr meanwhile
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ["ROLE_NAME"] = "dlrover-trainer"  # the saver is hosted in-process by local rank 0
    os.environ.setdefault("TORCHELASTIC_RUN_ID",
                          "fcpbench_%s" % os.environ.get("MASTER_PORT", os.getpid()))
    os.environ["LOCAL_WORLD_SIZE"] = str(world)
    os.environ["LOCAL_RANK"] = str(local_rank)
    os.environ["NODE_NUM"] = "1"
    os.environ["NODE_RANK"] = "0"
    os.environ.setdefault("DLROVER_LOG_LEVEL", "WARNING")

    if args.impl == "reference":
        run_reference_arm(args, rank, local_rank, world)
        return
    run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
