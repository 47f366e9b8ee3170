"""Megatron-LM flash checkpoint with the distributed optimizer (reference:
dlrover/trainer/torch/flash_checkpoint/megatron_dist_ckpt.py:150-709).

With ``--use-distributed-optimizer`` stock Megatron gathers every data-parallel
rank's optimizer shard on dp-rank 0 before saving.  Here every rank saves its
own ``{param, exp_avg, exp_avg_sq}`` shards — no gather — as the
``optim_states`` entry of its shard (file ``iter_XXXXXXX/rank_RRRRR/
distrib_optim.pt`` if it is ever persisted), and restores them in place.

The pieces on the data path are implemented here; everything that only
re-checks Megatron arguments defers to Megatron itself, which is imported
lazily (it is not a dependency of this package).
"""
import os
import sys

import torch
import torch.distributed as dist

from ..common.constants import CheckpointConstant
from ..common.log import default_logger as logger
from ..common.singleton import Singleton
from ..common.storage import (  # noqa: F401  (re-exported, as the reference does)
    CheckpointDeletionStrategy,
    KeepLatestStepStrategy,
    KeepStepIntervalStrategy,
    PosixDiskStorage,
    get_checkpoint_storage,
)
from ._hijack import _read_file
from .checkpointer import StorageType
from .megatron_engine import MegatronCheckpointEngine, MegatronDistCheckpointEngine


def _megatron():
    """The handful of Megatron symbols the adapter needs."""
    import importlib
    ns = {}
    try:
        training = importlib.import_module("megatron.training")
        ckpt = importlib.import_module("megatron.training.checkpointing")
        utils = importlib.import_module("megatron.training.utils")
    except ImportError:
        training = importlib.import_module("megatron")
        ckpt = importlib.import_module("megatron.checkpointing")
        utils = importlib.import_module("megatron.utils")
    core = importlib.import_module("megatron.core")
    ns["get_args"] = training.get_args
    ns["mpu"] = core.mpu
    ns["tensor_parallel"] = core.tensor_parallel
    ns["unwrap_model"] = utils.unwrap_model
    ns["print_rank_0"] = utils.print_rank_0
    for name in ("get_checkpoint_name", "get_rng_state", "get_checkpoint_tracker_filename",
                 "read_metadata", "find_checkpoint_rank_0", "check_checkpoint_args",
                 "set_checkpoint_version", "get_checkpoint_version", "fix_query_key_value_ordering",
                 "update_num_microbatches"):
        ns[name] = getattr(ckpt, name, None)
    try:
        ns["ChainedOptimizer"] = importlib.import_module("megatron.core.optimizer").ChainedOptimizer
    except (ImportError, AttributeError):
        ns["ChainedOptimizer"] = ()
    return ns


class MegatronDistCheckpointer(Singleton):
    def __init__(self, checkpoint_dir, storage=None, comm_backend="",
                 use_distributed_optimizer=False, save_timeout=CheckpointConstant.SAVE_TIMEOUT):
        self.storage = storage if storage else PosixDiskStorage()
        engine_cls = (MegatronDistCheckpointEngine if use_distributed_optimizer
                      else MegatronCheckpointEngine)
        self.engine = engine_cls(checkpoint_dir=checkpoint_dir, storage=self.storage,
                                 comm_backend=comm_backend, save_timeout=save_timeout)


def get_dist_optimizer_checkpoint_name(checkpoints_path, iteration, release=False):
    directory = "release" if release else "iter_{:07d}".format(iteration)
    rank = dist.get_rank() if dist.is_initialized() else 0
    return os.path.join(checkpoints_path, directory, "rank_%05d" % rank, "distrib_optim.pt")


def _param_state_slots(dist_optimizer):
    """Yield (bucket_idx, group_index, group_order, tensors) for every
    parameter shard this rank's distributed optimizer owns; `tensors` maps
    "param" and the inner optimizer's state names to the live tensors."""
    for gbuf_range_maps in dist_optimizer.gbuf_ranges:
        assert len(gbuf_range_maps) == 1, "single dtype supported, for now."
        for _, per_bucket in gbuf_range_maps.items():
            for bucket_idx, gbuf_range_map in enumerate(per_bucket):
                for model_param in gbuf_range_map["param_map"]:
                    group_index, group_order = \
                        dist_optimizer.model_param_group_index_map[model_param]
                    main_param = dist_optimizer.optimizer.param_groups[group_index][
                        "params"][group_order]
                    tensors = {"param": main_param,
                               **dist_optimizer.optimizer.state[main_param]}
                    yield bucket_idx, group_index, group_order, tensors


def get_parameter_state(dist_optimizer):
    """This rank's shard of the distributed optimizer as a nested dict
    ``{bucket: {group: {order: {"param", "exp_avg", "exp_avg_sq", ...}}}}`` of
    the LIVE tensors — nothing is copied or gathered; the pack kernel reads
    them where they are."""
    state = {}
    for bucket_idx, group_index, group_order, tensors in _param_state_slots(dist_optimizer):
        state.setdefault(bucket_idx, {}).setdefault(group_index, {})[group_order] = tensors
    return state


def get_chained_optimizer_parameter_state(chained_optimizer):
    return [get_parameter_state(o) if hasattr(o, "get_parameter_state") else None
            for o in chained_optimizer.chained_optimizers]


def load_parameter_state_from_state_dict(dist_optimizer, state_dict):
    """Copy restored shards into the live tensors (megatron_dist_ckpt.py:657-698
    in the reference).  On a GPU job `state_dict` already lives on the device
    (see _load_checkpoint_from_memory), so these are device-to-device copies;
    a shard restored in place (``engine.load(into=...)``) is skipped."""
    for bucket_idx, group_index, group_order, tensors in _param_state_slots(dist_optimizer):
        restored = state_dict[bucket_idx][group_index][group_order]
        for key, t in tensors.items():
            if restored[key] is not t:
                t.data.copy_(restored[key])


def load_chained_optimizer_parameter_state(chained_optimizer, states):
    for idx, optimizer in enumerate(chained_optimizer.chained_optimizers):
        if hasattr(optimizer, "load_parameter_state_from_state_dict"):
            load_parameter_state_from_state_dict(optimizer, states[idx] if states else None)


def save_checkpoint(iteration, model, optimizer, opt_param_scheduler,
                    num_floating_point_operations_so_far, storage_type=StorageType.DISK,
                    comm_backend="", deletion_strategy=None,
                    save_timeout=CheckpointConstant.SAVE_TIMEOUT):
    """Drop-in for Megatron's save_checkpoint.

    Args:
        deletion_strategy: optional CheckpointDeletionStrategy.
        save_timeout (int): seconds node 0 waits for all shards to persist.
    """
    mg = _megatron()
    args = mg["get_args"]()
    storage = get_checkpoint_storage(deletion_strategy)
    checkpointer = MegatronDistCheckpointer.singleton_instance(
        args.save, storage=storage, comm_backend=comm_backend,
        use_distributed_optimizer=args.use_distributed_optimizer, save_timeout=save_timeout)
    model = mg["unwrap_model"](model)
    mg["print_rank_0"]("saving checkpoint at iteration {:7d} to {}".format(iteration, args.save))
    rng_state = mg["get_rng_state"]()
    checkpoint_name = mg["get_checkpoint_name"](args.save, iteration)
    optim_checkpoint_name = get_dist_optimizer_checkpoint_name(args.save, iteration)

    dist_opter_state = {}
    if args.use_distributed_optimizer and not args.no_save_optim and optimizer is not None:
        if mg["ChainedOptimizer"] and isinstance(optimizer, mg["ChainedOptimizer"]):
            dist_opter_state = get_chained_optimizer_parameter_state(optimizer)
        else:
            dist_opter_state = get_parameter_state(optimizer)

    model_state_dict = {}
    mpu = mg["mpu"]
    if not dist.is_initialized() or mpu.get_data_modulo_expert_parallel_rank() == 0:
        model_state_dict["args"] = args
        model_state_dict["checkpoint_version"] = 3.0
        model_state_dict["iteration"] = iteration
        model_state_dict["num_floating_point_operations_so_far"] = \
            num_floating_point_operations_so_far
        if len(model) == 1:
            model_state_dict["model"] = model[0].state_dict_for_save_checkpoint()
        else:
            for i in range(len(model)):
                mpu.set_virtual_pipeline_model_parallel_rank(i)
                model_state_dict["model%d" % i] = model[i].state_dict_for_save_checkpoint()
        if not args.no_save_optim:
            if optimizer is not None:
                model_state_dict["optimizer"] = optimizer.state_dict()
            if opt_param_scheduler is not None:
                model_state_dict["opt_param_scheduler"] = opt_param_scheduler.state_dict()
        if not args.no_save_rng:
            model_state_dict["rng_state"] = rng_state

    ckpt_sds, paths = {}, {}
    if model_state_dict:
        ckpt_sds[CheckpointConstant.MODEL_STATES_NAME] = model_state_dict
        paths[CheckpointConstant.MODEL_STATES_NAME] = checkpoint_name
    if dist_opter_state:
        ckpt_sds[CheckpointConstant.OPTIM_STATES_NAME] = dist_opter_state
        paths[CheckpointConstant.OPTIM_STATES_NAME] = optim_checkpoint_name
    if storage_type == StorageType.MEMORY:
        ok = checkpointer.engine.save_to_memory(iteration, ckpt_sds, paths)
    else:
        ok = checkpointer.engine.save_to_storage(iteration, ckpt_sds, paths)
    if dist.is_initialized():
        dist.barrier()
    return ok


def _load_checkpoint_from_memory(checkpointer, into=None):
    # GPU jobs: fresh tensors on the current CUDA device through one verified
    # scatter launch; model.load_state_dict / load_parameter_state_from_state_dict
    # below then copy device-to-device instead of one pageable H2D per tensor
    step, state_dict = checkpointer.engine.load(
        into=into, device=checkpointer.engine.restore_device())
    return (state_dict.get(CheckpointConstant.MODEL_STATES_NAME, {}),
            state_dict.get(CheckpointConstant.OPTIM_STATES_NAME, {}),
            "iter_{:07d}".format(step), False)


def _load_base_checkpoint(load_dir, rank0=False):
    """(model_state_dict, opt_state_dict, checkpoint_name, release) from
    storage, following Megatron's tracker file."""
    mg = _megatron()
    tracker = mg["get_checkpoint_tracker_filename"](load_dir)
    if not os.path.isfile(tracker):
        if not rank0:
            mg["print_rank_0"]("WARNING: could not find the metadata file %s; will not load "
                               "any checkpoints and will start from random" % tracker)
        return None, None, "", False
    iteration, release = mg["read_metadata"](tracker)
    if rank0:
        checkpoint_name = mg["find_checkpoint_rank_0"](load_dir, iteration, release)
    else:
        checkpoint_name = mg["get_checkpoint_name"](load_dir, iteration, release)
    opt_name = get_dist_optimizer_checkpoint_name(load_dir, iteration, release)
    try:
        model_sd = _read_file(checkpoint_name, "cpu")
        opt_sd = _read_file(opt_name, "cpu") \
            if os.path.exists(opt_name) else {}
    except BaseException as e:  # noqa: BLE001 - Megatron exits here too
        mg["print_rank_0"]("could not load the checkpoint: %s" % e)
        sys.exit()
    return model_sd, opt_sd, checkpoint_name, release


def load_checkpoint(model, optimizer, opt_param_scheduler, load_arg="load", strict=True,
                    comm_backend="", deletion_strategy=None):
    """Drop-in for Megatron's load_checkpoint: shared memory first, storage
    second.  Returns ``(iteration, num_floating_point_operations_so_far)``."""
    mg = _megatron()
    args = mg["get_args"]()
    load_dir = getattr(args, load_arg)
    storage = get_checkpoint_storage(deletion_strategy)
    checkpointer = MegatronDistCheckpointer.singleton_instance(
        args.save, storage=storage, comm_backend=comm_backend,
        use_distributed_optimizer=args.use_distributed_optimizer)
    model = mg["unwrap_model"](model)
    model_sd, opt_sd, checkpoint_name, release = _load_checkpoint_from_memory(checkpointer)
    if not model_sd:
        model_sd, opt_sd, checkpoint_name, release = _load_base_checkpoint(load_dir)
    if model_sd is None:
        return 0, 0
    if mg["set_checkpoint_version"]:
        mg["set_checkpoint_version"](model_sd.get("checkpoint_version", 0))
    if args.finetune or release:
        iteration = 0
    else:
        iteration = model_sd.get("iteration", model_sd.get("total_iters", 0))
    flops = model_sd.get("num_floating_point_operations_so_far", 0)
    if "args" in model_sd and not args.finetune and mg["check_checkpoint_args"]:
        mg["check_checkpoint_args"](model_sd["args"])
        args.consumed_train_samples = getattr(model_sd["args"], "consumed_train_samples", 0)
        if mg["update_num_microbatches"]:
            mg["update_num_microbatches"](consumed_samples=args.consumed_train_samples)
        args.consumed_valid_samples = getattr(model_sd["args"], "consumed_valid_samples", 0)
    if len(model) == 1:
        model[0].load_state_dict(model_sd["model"], strict=strict)
    else:
        for i in range(len(model)):
            mg["mpu"].set_virtual_pipeline_model_parallel_rank(i)
            model[i].load_state_dict(model_sd["model%d" % i], strict=strict)
    if mg["get_checkpoint_version"] and mg["fix_query_key_value_ordering"]:
        mg["fix_query_key_value_ordering"](model, mg["get_checkpoint_version"]())
    if not release and not args.finetune and not args.no_load_optim:
        try:
            if optimizer is not None:
                optimizer.load_state_dict(model_sd["optimizer"])
            if opt_sd and optimizer is not None:
                if mg["ChainedOptimizer"] and isinstance(optimizer, mg["ChainedOptimizer"]):
                    load_chained_optimizer_parameter_state(optimizer, opt_sd)
                else:
                    load_parameter_state_from_state_dict(optimizer, opt_sd)
            if opt_param_scheduler is not None:
                opt_param_scheduler.load_state_dict(
                    model_sd.get("opt_param_scheduler", model_sd.get("lr_scheduler")))
        except KeyError:
            mg["print_rank_0"]("Unable to load optimizer from checkpoint %s. Specify "
                               "--no-load-optim or --finetune to prevent attempting to load "
                               "the optimizer state, exiting ..." % checkpoint_name)
            sys.exit()
    if not release and not args.finetune and not args.no_load_rng:
        try:
            rng = model_sd["rng_state"]
            rng = rng[mg["mpu"].get_data_parallel_rank()] if args.data_parallel_random_init \
                else rng[0]
            import random

            import numpy as np
            random.setstate(rng["random_rng_state"])
            np.random.set_state(rng["np_rng_state"])
            torch.set_rng_state(rng["torch_rng_state"])
            torch.cuda.set_rng_state(rng["cuda_rng_state"])
            mg["tensor_parallel"].get_cuda_rng_tracker().set_states(rng["rng_tracker_states"])
        except KeyError:
            mg["print_rank_0"]("Unable to load rng state from checkpoint %s" % checkpoint_name)
            sys.exit()
    if dist.is_initialized():
        dist.barrier()
    logger.info("loaded checkpoint %s at iteration %s", checkpoint_name, iteration)
    return iteration, flops
