"""TEST INFRASTRUCTURE.  Generates tests/golden/ref_fsdp_writer.pt by running
the REFERENCE's own FSDP shm writer —
dlrover/trainer/torch/flash_checkpoint/fsdp_engine.py:85-221
(_write_memory_from_list, _write_item, SharedMemoryWriter) — through
``torch.distributed.checkpoint.save`` on the state of oracle/fsdp_case.py.

Recipe of SURVEY §8(c): the reference's writer does not instantiate on torch
2.11 (two new abstract methods), so a 2-method subclass shim is used, nothing
in the reference is edited; the planner is ``DefaultSavePlanner(
dedup_save_to_lowest_rank=True)`` (the torch-2.0 behaviour the reference was
written against); one rank, gloo, CPU.

Run in the build container only (needs /root/reference or baseline/_ref):

    python oracle/gen_golden_fsdp.py

The fixture records: the inputs, the file name, the segment size the reference
allocated, the bytes of the used prefix, every item in plan order with its
_StorageInfo(relative_path, offset, length), the non-sharded data the writer
puts in the meta, and the DCP Metadata (storage_data as plain tuples, so that
loading the fixture needs no reference class).
"""
import io
import os
import sys

import torch
import torch.distributed as dist
import torch.distributed.checkpoint as dist_cp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
OUT = os.path.join(HERE, "..", "tests", "golden", "ref_fsdp_writer.pt")


def main():
    from oracle import fsdp_case, reference_closure
    ref = reference_closure.import_reference("fcp_golden_fsdp_%d" % os.getpid())
    from dlrover.trainer.torch.flash_checkpoint import fsdp_engine as ref_fsdp

    class Writer(ref_fsdp.SharedMemoryWriter):  # the 2-method shim
        def reset(self, checkpoint_id=None):
            pass

        @classmethod
        def validate_checkpoint_id(cls, checkpoint_id):
            return True

    dist.init_process_group("gloo", rank=0, world_size=1,
                            init_method="tcp://127.0.0.1:29541")
    inputs = fsdp_case.make_inputs()
    sd = fsdp_case.build(inputs, "cpu")
    handler = ref.SharedMemoryHandler(61, host=True)
    writer = Writer(handler)
    dist_cp.save(sd, storage_writer=writer,
                 planner=dist_cp.DefaultSavePlanner(dedup_save_to_lowest_rank=True))
    md = writer.metadata["dcp_metadata"]
    items, used = [], 0
    for index, info in md.storage_data.items():  # insertion order == plan order
        items.append({"fqn": index.fqn,
                      "offset": None if index.offset is None else tuple(index.offset),
                      "index": index.index, "relative_path": info.relative_path,
                      "at": info.offset, "length": info.length})
        used = max(used, info.offset + info.length)
    no_shard = {}
    for fqn, data in writer.metadata["no_shard_data"].items():
        if isinstance(data, io.BytesIO):
            no_shard[fqn] = ("bytes", data.getvalue())
        else:
            no_shard[fqn] = ("tensor", data.clone())
    plain_md = dist_cp.metadata.Metadata(
        state_dict_metadata=md.state_dict_metadata,
        planner_data=md.planner_data,
        storage_data={i: (v.relative_path, v.offset, v.length)
                      for i, v in md.storage_data.items()})
    fixture = {"torch": torch.__version__, "inputs": inputs,
               "file_name": writer.file_name, "shm_size": handler.shared_memory.size,
               "used": used, "prefix": bytes(handler.shared_memory.buf[:used]),
               "items": items, "no_shard": no_shard, "dcp_metadata": plain_md}
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    torch.save(fixture, OUT)
    print("file", writer.file_name, "allocated", handler.shared_memory.size, "used", used,
          "items", len(items))
    print("wrote", os.path.normpath(OUT), os.path.getsize(OUT), "bytes")
    handler.unlink()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
