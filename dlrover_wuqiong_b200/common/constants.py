"""Names shared between trainer and agent on the flash-checkpoint path.
Mirrors dlrover/python/common/constants.py:205-231,294-298 (only the members
the path reads)."""


class NodeEnv(object):
    NODE_ID = "NODE_ID"
    NODE_NUM = "NODE_NUM"
    NODE_RANK = "NODE_RANK"
    NODE_TYPE = "NODE_TYPE"
    WORKER_RANK = "WORKER_RANK"  # pre-0.3.0 spelling of NODE_RANK
    TORCHELASTIC_RUN_ID = "TORCHELASTIC_RUN_ID"


class CheckpointConstant(object):
    TRACER_FILE_NAME = "dlrover_latest.txt"
    MODEL_STATES_NAME = "model_states"
    OPTIM_STATES_NAME = "optim_states"
    SAVE_TIMEOUT = 600
