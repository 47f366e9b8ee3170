"""Flash-checkpoint bench: one "step" = one save of a rank's shard from HBM into
host shared memory (durable, CRC-32C per 2 MiB chunk) + one restore of it back
into the live tensors (CRC verified).  See DESIGN.md §7 for the definitions.

    python bench.py --gpus 1 --steps 5 --warmup 3
    torchrun --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference's CPU path (oracle port)

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
NCU_TRAFFIC_BYTES = {("gpt2-xl-1.5B-adamw-fp32", 0): 39.10e9}
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


def build_sharded_state(workload, device, seed):
    """Per-rank shard of configs 3 / 4 as the checkpointers hand it to the
    engine: low-precision model shard + fp32 optimizer state (+ CPU rng
    states for the Megatron one)."""
    c = SHARDED[workload]
    lowp = torch.bfloat16 if c["kind"] == "llama" else torch.float16
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    model, state = {}, {}
    for idx, (name, numel) in enumerate(sharded_shapes(workload)):
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
        return build_sharded_state(workload, device, seed)
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


# --------------------------------------------------------- CPU baseline legs
def cpu_reference_steps(workload, steps, warmup, sample_bytes):
    """The reference's CPU path (oracle port, SAVER:174-206 + 112-137) on a
    bounded prefix of the workload held in host memory: per step one save
    into a shared mapping + one restore into the tensors.  Returns
    (GB/s, seconds per step, sample description, threads)."""
    from oracle import oracle
    sd = build_state(workload, "cpu", 1234, max_bytes=sample_bytes)
    meta, total = oracle.layout(sd)
    mm = mmap.mmap(-1, total, flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)
    into = oracle.traverse(sd, lambda v: torch.empty_like(v) if torch.is_tensor(v) else v)

    def step():
        oracle.ref_save_into(sd, meta, mm)
        oracle.ref_load_from(meta, mm, into)

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    ok = torch.equal(into["model"]["transformer.wte.weight"], sd["model"]["transformer.wte.weight"])
    assert ok, "CPU baseline round trip failed"
    n_tensors = len(oracle.flatten(sd)[0])
    sample = "%s prefix: %d tensors, %.2f GB, save+restore" % (workload, n_tensors, total / GB)
    del into, sd
    mm.close()
    return 2 * total / dt / GB, dt, sample, torch.get_num_threads()


def reference_loop_on_gpu_tensors(sd, sample_bytes):
    """SAVER:174-206 / 112-137 (oracle port) with the tensors on the GPU: one
    synchronous pageable D2H per tensor on save, one pageable H2D per tensor
    on restore."""
    from oracle import oracle
    model, state, used = {}, {}, 0
    for i, (name, w) in enumerate(sd["model"].items()):
        if i not in sd["optimizer"]["state"]:
            break
        if used + w.numel() * 12 > sample_bytes and model:
            break
        model[name] = w
        state[i] = sd["optimizer"]["state"][i]
        used += w.numel() * 12 + 4
    sample = {"model": model, "optimizer": {"state": state}}
    meta, total = oracle.layout(sample)
    mm = mmap.mmap(-1, total, flags=mmap.MAP_SHARED | mmap.MAP_POPULATE)
    times = []
    for _ in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        oracle.ref_save_into(sample, meta, mm)
        oracle.ref_load_from(meta, mm, sample)
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
    mm.close()
    dt = min(times[1:])
    return {"value": 2 * total / dt / GB, "unit": "GB/s", "s_per_step": dt,
            "sample": "prefix of the bench state: %d tensors, %.2f GB, save+restore"
                      % (len(oracle.flatten(sample)[0]), total / GB),
            "what": "reference per-tensor torch copy_ loop, CUDA tensors <-> pageable mapping"}


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


def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:  # noqa: BLE001
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path
    (oracle port) on the host cores; rank 0 only."""
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    gbs, dt, sample, threads = cpu_reference_steps(
        args.workload, args.steps, args.warmup, int(args.cpu_sample_gb * GB))
    line = {
        "impl": "reference", "metric": "flash_ckpt_save_plus_restore_GBps", "value": gbs,
        "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": args.workload, "sample": sample,
                   "what": "reference CPU path: per-tensor torch copy_ into a shared "
                           "mapping + per-tensor copy_ back (oracle port of "
                           "ckpt_saver.py:174-206,112-137)"},
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "host": host_info(),
    }
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)


def run_ours(args, rank, local_rank, world):
    from dlrover_wuqiong_b200 import _native
    from dlrover_wuqiong_b200.ckpt_saver import (
        DLROVER_CKPT_CONFIG_KEY, FCP_CHECKSUM_KEY, AsyncCheckpointSaver, CheckpointConfig)
    from dlrover_wuqiong_b200.common.multi_process import clear_sock_dir
    from dlrover_wuqiong_b200.flash_checkpoint.ddp import DdpCheckpointer, StorageType

    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    # ---- size guards (a box driven out of memory is a strike)
    hi = host_info()
    workload = args.workload
    B = workload_bytes(workload)
    avail = hi.get("mem_available_gb", 1e9) * GB
    need_host = world * B * 1.15 + (2.5 * args.cpu_sample_gb * GB if world == 1 else 0)
    free_hbm, _ = torch.cuda.mem_get_info(device)
    note = None
    try:
        st = os.statvfs("/dev/shm")
        shm_free = st.f_bavail * st.f_frsize
    except OSError:
        shm_free = avail
    if need_host > 0.85 * avail or world * B * 1.05 > shm_free or 2.3 * B > free_hbm:
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

    # ---- parity at full size: sampled chunks of the shm image against the
    # oracle's CRC-32C, and restored tensors against a snapshot
    meta = handler.metadata.get(local=True)
    sums = np.asarray(meta[FCP_CHECKSUM_KEY]["sums"], dtype=np.uint32)
    chunk = meta[FCP_CHECKSUM_KEY]["chunk_bytes"]
    parity = {"chunks_checked": 0, "tensors_checked": 0}
    if rank == 0:
        from oracle import oracle
        shm = np.frombuffer(handler.shared_memory.buf, dtype=np.uint8)
        for c in sorted({0, 1, len(sums) // 3, len(sums) // 2, len(sums) - 2, len(sums) - 1}):
            if 0 <= c < len(sums):
                assert oracle.crc32c(shm[c * chunk:(c + 1) * chunk]) == int(sums[c]), \
                    "chunk %d CRC differs from oracle" % c
                parity["chunks_checked"] += 1
        del shm
    names = list(sd["model"].keys())
    snap_names = [names[1], names[len(names) // 2], names[-2]]
    snaps = {n: sd["model"][n].clone() for n in snap_names}
    avg1 = sd["optimizer"]["state"][1]["exp_avg"].clone()
    for n in snaps:
        sd["model"][n].zero_()
    sd["optimizer"]["state"][1]["exp_avg"].add_(1.0)
    ckpt.load_checkpoint(into=sd)
    for n, s in snaps.items():
        assert torch.equal(sd["model"][n], s), "restore mismatch in %s" % n
        parity["tensors_checked"] += 1
    assert torch.equal(sd["optimizer"]["state"][1]["exp_avg"], avg1)
    parity["tensors_checked"] += 1
    del snaps, avg1

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
    for _ in range(K):
        plan.save_async(staging_ptr, addr, mode, _native.SUM_CRC32C, stream.cuda_stream)
        step_sums = plan.save_wait()
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
    prof = None
    if os.getenv("FCP_PROFILE_BLOCKING") and rank == 0:
        import cProfile
        prof = cProfile.Profile()
    for k in range(K):
        wrapped = {"model_states": sd,
                   DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=200 + k, paths={"model_states": path})}
        torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        if prof:
            prof.enable()
        pending = handler.save_state_dict_async(wrapped)
        pending.wait_packed()
        if prof:
            prof.disable()
        block_s.append(time.perf_counter() - t0)
        pending.wait()
    if prof:
        import pstats
        pstats.Stats(prof, stream=sys.stderr).sort_stats("cumulative").print_stats(18)
    clocks = sampler.stop() if rank == 0 else None

    # ---- leg 4: cold restore — what a restarted trainer sees: attach to the
    # existing segment (not pinned in this "new" process) and restore
    from dlrover_wuqiong_b200.ckpt_saver import SharedMemoryHandler
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

    # the true "before" on GPU-resident state: the reference's per-tensor
    # pageable copy_ loop (oracle port) fed with CUDA tensors, on a prefix of
    # the very same state dict
    ref_gpu = None
    gpt2 = workload in WORKLOADS
    if rank == 0 and world == 1 and not args.no_cpu_baseline and gpt2:
        ref_gpu = reference_loop_on_gpu_tensors(sd, int(args.cpu_sample_gb * GB))

    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and gpt2:
        gbs, dt, sample, threads = cpu_reference_steps(workload, 3, 1, int(args.cpu_sample_gb * GB))
        cpu_base = {"value": gbs, "unit": "GB/s", "cores": threads, "kind": "port",
                    "sample": sample, "s_per_step": dt}

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
                                         "pack launches, ncu --set full, "
                                         "profiles/r01_ncu_pack_v3_summary.md)",
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch_set": 2 * B},
            "cpu_baseline": cpu_base,
            "reference_loop_gpu_source": ref_gpu,
            "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": B,
                    "d2h_bytes_per_step": B, "s_per_step": e2e_s / K,
                    "api": "DdpCheckpointer.save_checkpoint(MEMORY) + load_checkpoint(into=)"},
            "gpu_launches": launches,
            "clocks": clocks, "parity": parity, "host": hi,
        }
        print(json.dumps(line), file=_REAL_STDOUT, flush=True)

    ckpt.close()
    if dist.is_initialized():
        dist.barrier()
    if local_rank == 0:
        saver = AsyncCheckpointSaver.get_ckpt_saver()
        if saver is not None:
            saver.close()
        clear_sock_dir()
    if dist.is_initialized():
        dist.destroy_process_group()


def _iter_tensors(v):
    if isinstance(v, dict):
        for x in v.values():
            yield from _iter_tensors(x)
    elif isinstance(v, list):
        for x in v:
            yield from _iter_tensors(x)
    elif torch.is_tensor(v):
        yield v


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="gpt2-xl-1.5B-adamw-fp32", choices=sorted(WORKLOADS) + sorted(SHARDED))
    ap.add_argument("--cpu-sample-gb", type=float, default=4.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
        run_reference_arm(args, rank, world)
        return
    run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
