"""Logger of the package (role of dlrover/python/common/log.py:35-45)."""
import logging
import os
import sys

_FMT = "[%(asctime)s] [%(levelname)s] [%(filename)s:%(lineno)d:%(funcName)s] %(message)s"


def _make():
    lg = logging.getLogger("dlrover_wuqiong_b200")
    if not lg.handlers:
        h = logging.StreamHandler(sys.stderr)
        h.setFormatter(logging.Formatter(_FMT))
        lg.addHandler(h)
        lg.propagate = False
    lg.setLevel(os.getenv("DLROVER_LOG_LEVEL", "INFO").upper())
    return lg


default_logger = _make()
