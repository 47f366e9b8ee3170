"""ClassMeta: how the trainer tells the agent which saver / storage class to
instantiate (dlrover/python/common/serialize.py:38-42)."""
from dataclasses import dataclass, field
from typing import Dict


@dataclass
class ClassMeta:
    module_path: str = ""
    class_name: str = ""
    kwargs: Dict[str, object] = field(default_factory=dict)
