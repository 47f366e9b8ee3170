"""Build libfcp.so in-tree for sm_100a (nvcc cross-compiles without a GPU).

    python -m dlrover_wuqiong_b200.csrc.build [--force]
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SOURCES = ["fcp.cu", "fcp_host.cpp"]
HEADERS = ["fcp_crc.h", "fcp_kernels.cuh", "fcp_kernels2.cuh",
           os.path.join("..", "..", "include", "fcp.h")]
TARGET = os.path.join(HERE, "libfcp.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3",
    "-std=c++17", "-shared", "-Xcompiler", "-fPIC",
]  # static cudart (nvcc default): self-contained, independent of torch's copy


def _stale():
    if not os.path.exists(TARGET):
        return True
    t = os.path.getmtime(TARGET)
    return any(os.path.getmtime(os.path.join(HERE, f)) > t
               for f in SOURCES + HEADERS)


def build(force=False, verbose=False):
    if not force and not _stale():
        return TARGET
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
        ["-o", TARGET] + [os.path.join(HERE, s) for s in SOURCES]
    subprocess.check_call(cmd, cwd=HERE)
    return TARGET


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
