"""In-tree build of libddim_b200.so (hand-written sm_100a kernels + the C ABI of include/ddim_b200.h).

nvcc cross-compiles without a GPU; the .so is git-ignored but travels to the GPU box with the snapshot.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libddim_b200.so")
SOURCES = ["runtime.cu", "gemm_host.cu", "elementwise.cu", "train_host.cu", "train_elementwise.cu", "conv_boundary.cu", "denoise2d.cu", "ema.cu", "optim.cu", "pfft2wav.cu"]
HEADERS = ["common.cuh", "umma_gemm.cuh", "conv_wide.cuh", "conv_window.cuh", "conv_out_tc.cuh", "conv_down_wide.cuh", "conv_up_wide.cuh", "gemm_tn.cuh", "wgrad_wide.cuh", "wgrad_updown_wide.cuh", "tmap.cuh", os.path.join("..", "..", "include", "ddim_b200.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O2", "-cudart", "static", "-Xptxas", "-v",
]
# --use_fast_math (approximate division / sqrt, flush-to-zero) only for the translation units whose arithmetic is bf16-storage GEMM
# epilogues and activation passes; the fp32 kernels that promise the reference's operation order or bit-exactness (DDIM update,
# q-sample, loss, optimizers, EMA, denoise_2d, the WAV codec) are compiled with IEEE division / sqrt and denormals
FAST_MATH = {"gemm_host.cu", "train_host.cu", "train_elementwise.cu", "conv_boundary.cu"}


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _digest() -> str:
    h = hashlib.sha256()
    for f in SOURCES + HEADERS:
        with open(os.path.join(CSRC, f), "rb") as fh:
            h.update(fh.read())
    h.update((" ".join(NVCC_FLAGS) + " fast:" + ",".join(sorted(FAST_MATH))).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(LIBDIR, exist_ok=True)
    stamp = os.path.join(LIBDIR, "libddim_b200.sha256")
    digest = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read().strip() == digest:
        return LIB
    objs = []
    log = []
    for src in SOURCES:
        obj = os.path.join(LIBDIR, src.replace(".cu", ".o"))
        cmd = [_nvcc(), *NVCC_FLAGS, *(["--use_fast_math"] if src in FAST_MATH else []), "-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log.append(r.stderr)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError(f"nvcc failed on {src}")
        objs.append(obj)
    cmd = [_nvcc(), "-shared", "-cudart", "static", "-o", LIB, *objs]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    with open(os.path.join(LIBDIR, "ptxas.log"), "w") as fh:
        fh.write("\n".join(log))
    with open(stamp, "w") as fh:
        fh.write(digest)
    if verbose:
        sys.stderr.write("\n".join(log))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
