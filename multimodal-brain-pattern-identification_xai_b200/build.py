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

SOURCES = ["abi.cu", "stft_image.cu", "melspec.cu", "melspec_tc.cu", "melspec_ts.cu", "eegtransform.cu", "spec_condition.cu", "classifier_eeg.cu", "classifier_spec.cu"]
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


def _sha(paths, extra=b"") -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode() + extra)
    for path in paths:
        with open(path, "rb") as fh:
            h.update(os.path.basename(path).encode())
            h.update(fh.read())
    return h.hexdigest()


def _headers():
    return [os.path.join(CSRC, n) for n in sorted(os.listdir(CSRC)) if n.endswith((".h", ".cuh"))] + \
        [os.path.join(INCLUDE, "eegspec.h")]


def _defines():
    """Extra -D flags for debug builds (tools only), e.g. EEGSPEC_BUILD_DEFINES=EEGSPEC_TIMELINE."""
    return ["-D" + d for d in os.environ.get("EEGSPEC_BUILD_DEFINES", "").split(",") if d]


def _fingerprint() -> str:
    return _sha([os.path.join(CSRC, s) for s in SOURCES] + _headers(), " ".join(_defines()).encode())


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile the library if the sources changed; return its path.

    Every translation unit is compiled to its own object (in parallel, cached by a hash of the
    source and all headers) and the objects are linked into the shared library."""
    from concurrent.futures import ThreadPoolExecutor

    os.makedirs(LIB_DIR, exist_ok=True)
    obj_dir = os.path.join(LIB_DIR, "obj")
    os.makedirs(obj_dir, exist_ok=True)
    fp = _fingerprint()
    if not force and os.path.exists(LIB_PATH) and os.path.exists(STAMP):
        with open(STAMP) as fh:
            if fh.read().strip() == fp:
                return LIB_PATH
    compile_flags = [f for f in NVCC_FLAGS if f not in ("-shared",)]
    hdrs = _headers()

    def compile_one(src: str):
        obj = os.path.join(obj_dir, src[:-3] + ".o")
        stamp = obj + ".stamp"
        want = _sha([os.path.join(CSRC, src)] + hdrs, " ".join(_defines()).encode())
        if not force and os.path.exists(obj) and os.path.exists(stamp):
            with open(stamp) as fh:
                if fh.read().strip() == want:
                    return obj, ""
        cmd = [_nvcc(), *compile_flags, *_defines(), "-I", INCLUDE, "-c"]
        if verbose:
            cmd += ["-Xptxas", "-v"]
        cmd += ["-o", obj, os.path.join(CSRC, src)]
        proc = subprocess.run(cmd, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + proc.stdout + proc.stderr)
        with open(stamp, "w") as fh:
            fh.write(want)
        return obj, proc.stderr

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as pool:
        results = list(pool.map(compile_one, SOURCES))
    if verbose:
        for _, err in results:
            sys.stderr.write(err)
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", LIB_PATH, *[obj for obj, _ in results]]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + " ".join(cmd) + "\n" + proc.stdout + proc.stderr)
    with open(STAMP, "w") as fh:
        fh.write(fp)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
