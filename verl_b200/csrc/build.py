"""Build libvpg_b200.so (hand-written sm_100a CUDA behind a C ABI) in-tree with nvcc.

Usage: python -m verl_b200.csrc.build [--force] [--verbose]
Output: verl_b200/lib/libvpg_b200.so  (git-ignored; travels to the GPU box with the repo snapshot)
"""

from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

CSRC = Path(__file__).resolve().parent
PKG = CSRC.parent
LIB_DIR = PKG / "lib"
BUILD_DIR = CSRC / "build"
LIB_PATH = LIB_DIR / "libvpg_b200.so"

SOURCES = ["common.cu", "lce.cu", "token_ops.cu", "adv_ops.cu", "rmpad.cu", "logits_ops.cu", "seqlen_balance.cu", "packed_seqs.cu", "cov_loss.cu"]
HEADERS = ["common.h", "ptx.cuh", "tc_gemm.cuh", "reduce.cuh", "token_common.cuh", "row_copy.cuh", "../../include/vpg_b200.h"]

NVCC_FLAGS = [
    "-gencode",
    "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-O3",
    "-std=c++17",
    "--use_fast_math",
    "-Xcompiler",
    "-fPIC",
]
# NOTE: --use_fast_math only affects the tcgen05 epilogues' plain float ops (ftz); every transcendental in the
# parity-critical token ops is written with explicit accurate intrinsics (see token_ops.cu).


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libvpg_b200.so cannot be built")


def _fingerprint(flags_per_src: dict[str, list[str]]) -> str:
    h = hashlib.sha256()
    for name in SOURCES + HEADERS:
        p = (CSRC / name).resolve()
        if p.exists():
            h.update(name.encode())
            h.update(p.read_bytes())
    h.update(repr(NVCC_FLAGS).encode())
    h.update(repr(sorted(flags_per_src.items())).encode())
    return h.hexdigest()


def _flags_for(src: str) -> list[str]:
    flags = list(NVCC_FLAGS)
    if src in ("token_ops.cu", "adv_ops.cu", "rmpad.cu", "logits_ops.cu", "packed_seqs.cu", "seqlen_balance.cu", "cov_loss.cu"):
        # bit-faithful fp32 arithmetic for the parity-critical memory-bound ops
        flags.remove("--use_fast_math")
        flags += ["--fmad=false"] if src in ("token_ops.cu", "adv_ops.cu", "cov_loss.cu") else []
    return flags


def build(force: bool = False, verbose: bool = False) -> Path:
    srcs = [s for s in SOURCES if (CSRC / s).exists()]
    flags_per_src = {s: _flags_for(s) for s in srcs}
    fp = _fingerprint(flags_per_src)
    stamp = BUILD_DIR / "fingerprint.txt"
    if not force and LIB_PATH.exists() and stamp.exists() and stamp.read_text() == fp:
        return LIB_PATH
    nvcc = _nvcc()
    BUILD_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)

    def compile_one(src: str) -> Path:
        obj = BUILD_DIR / (Path(src).stem + ".o")
        cmd = [nvcc, *flags_per_src[src], "-I", str(CSRC), "-c", str(CSRC / src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas")
            cmd.insert(2, "-v")
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr, flush=True)
        return obj

    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(compile_one, srcs))
    link = [nvcc, "-shared", "-o", str(LIB_PATH), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a",
            "-Xcompiler", "-fPIC"]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(fp)
    return LIB_PATH


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(p)
