"""Build ``lib/libeegspec.so`` in-tree with nvcc for sm_100a (no torch headers involved).

The library is a plain C-ABI shared object (``include/eegspec.h``); the CUDA runtime is linked
statically so that it loads (and exports its symbols) on a machine without a GPU driver.

    python -m b200spec.build        # or:  python __graft_entry__.py
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_DIR = os.path.join(PKG_DIR, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libeegspec.so")
STAMP = os.path.join(LIB_DIR, "libeegspec.stamp")
INCLUDE = os.path.join(os.path.dirname(PKG_DIR), "include")

SOURCES = ["abi.cu", "stft_image.cu", "melspec.cu", "melspec_tc.cu", "melspec_chain.cu", "melspec_roles.cu"]
NVCC_FLAGS = [
    "-O3", "-std=c++17",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-Xcompiler", "-fPIC",
    "-shared", "-cudart", "static",
]


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: the eegspec CUDA library cannot be built")
    return exe


def _fingerprint() -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    files = sorted(os.listdir(CSRC)) + [os.path.join(INCLUDE, "eegspec.h")]
    for name in files:
        path = name if os.path.isabs(name) else os.path.join(CSRC, name)
        with open(path, "rb") as fh:
            h.update(name.encode())
            h.update(fh.read())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile the library if the sources changed; return its path."""
    os.makedirs(LIB_DIR, exist_ok=True)
    fp = _fingerprint()
    if not force and os.path.exists(LIB_PATH) and os.path.exists(STAMP):
        with open(STAMP) as fh:
            if fh.read().strip() == fp:
                return LIB_PATH
    cmd = [_nvcc(), *NVCC_FLAGS, "-I", INCLUDE]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += ["-o", LIB_PATH, *[os.path.join(CSRC, s) for s in SOURCES]]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + proc.stdout + proc.stderr)
    if verbose:
        sys.stderr.write(proc.stderr)
    with open(STAMP, "w") as fh:
        fh.write(fp)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
