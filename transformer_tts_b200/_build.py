"""In-tree build of the sm_100a CUDA library (``csrc/libttsb200.so``).

nvcc cross-compiles without a GPU, so this runs on the CPU build box; the ``.so`` is git-ignored but
travels to the GPU box with the repo snapshot.  No JIT cache, no torch extension machinery: the
library has a plain C ABI (``include/ttsb200.h``) and is loaded with ctypes.
"""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

CSRC = Path(__file__).resolve().parent / "csrc"
LIB_PATH = CSRC / "libttsb200.so"
SOURCES = ["api.cu", "gemm.cu", "attention.cu", "rowwise.cu", "wgrad.cu", "train_rows.cu", "attention_bwd.cu", "precise.cu", "decode.cu", "decode_cluster.cu"]
HEADERS = [CSRC / "common.cuh", CSRC / "gemm_kernel.cuh", CSRC / "gemm_epilogue.inc", CSRC / "gemm_host.cuh", CSRC / "ffn_fused.cuh",
           CSRC.parent.parent / "include" / "ttsb200.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found: cannot build csrc/libttsb200.so")


def is_stale() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    deps = [CSRC / s for s in SOURCES] + HEADERS
    return any(d.stat().st_mtime > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every .cu for sm_100a and link the shared library.  Idempotent."""
    if not force and not is_stale():
        return LIB_PATH
    nvcc = _nvcc()
    objs = []
    bdir = CSRC / "build"
    bdir.mkdir(exist_ok=True)
    procs = []
    for s in SOURCES:
        obj = bdir / (s[:-3] + ".o")
        # TTSB200_NVCC_FLAGS: extra -D switches for experiment builds (tools/exp), empty in a product build
        cmd = [nvcc, *NVCC_FLAGS, *os.environ.get("TTSB200_NVCC_FLAGS", "").split(), "-c", str(CSRC / s), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(str(obj))
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {s}:\n{out}")
        if verbose and out:
            print(out)
    cmd = [nvcc, "-shared", "-o", str(LIB_PATH), *objs, "-cudart", "static"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}")
    return LIB_PATH


if __name__ == "__main__":
    import sys

    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
