"""TEST INFRASTRUCTURE — snapshot the reference's own hot-path modules and their tests into ``oracle/_ref/``.

    python -m oracle.vendor_ref            (build container only: needs /root/reference)

The reference is pure Python, so "building ``oracle/_ref``" is a file snapshot: the files below are copied UNMODIFIED,
with their relative paths, into ``oracle/_ref/`` (git-ignored — the repository history never holds reference sources —
but NOT gpurun-ignored, so the snapshot travels to the GPU box like a built ``.so``).  ``MANIFEST.json`` records the
sha256 of every file.  Nothing under ``verl_b200/`` imports it; its users are

* ``bench.py --impl reference`` / ``cpu_baseline``: the reference's ``FusedLinearForPPO`` + ``core_algos`` on the host
  cores (``kind: "reference"``);
* ``scripts/bench_ref_gpu.py``: the reference's Triton op and ``FusedLinearForPPO`` timed on the same B200;
* ``tests/test_vendored_gpu.py``: the reference's own acceptance tests pointed at ``verl_b200`` through
  ``compat.install()``.
"""

from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEST = os.path.join(ROOT, "oracle", "_ref")
SRC = os.environ.get("VERL_REFERENCE_SRC", "/root/reference")

FILES = [
    # the path's modules
    "verl/utils/kernel/__init__.py",
    "verl/utils/kernel/kernels.py",
    "verl/utils/kernel/linear_cross_entropy.py",
    "verl/utils/experimental/__init__.py",
    "verl/utils/experimental/torch_functional.py",
    "verl/utils/torch_functional.py",
    "verl/utils/device.py",
    "verl/trainer/ppo/__init__.py",
    "verl/trainer/ppo/core_algos.py",
    "verl/trainer/ppo/metric_utils.py",
    "verl/utils/metric/__init__.py",
    "verl/utils/metric/utils.py",
    "verl/utils/import_utils.py",
    # the reference's own tests of the path
    "tests/utils/test_linear_cross_entropy.py",
    "tests/utils/test_linear_cross_entropy_tp.py",
    "tests/utils/test_torch_functional.py",
    "tests/trainer/ppo/test_core_algos_on_cpu.py",
    "tests/trainer/ppo/test_metric_utils_on_cpu.py",
]


def available() -> bool:
    return os.path.isfile(os.path.join(DEST, "MANIFEST.json"))


def vendor(force: bool = False) -> str:
    """Copy FILES from the reference tree into oracle/_ref (no-op when the manifest is current)."""
    if not os.path.isdir(os.path.join(SRC, "verl")):
        raise RuntimeError(f"reference tree not found at {SRC}")
    manifest = {}
    for rel in FILES:
        with open(os.path.join(SRC, rel), "rb") as f:
            manifest[rel] = hashlib.sha256(f.read()).hexdigest()
    mpath = os.path.join(DEST, "MANIFEST.json")
    if not force and os.path.isfile(mpath):
        try:
            if json.load(open(mpath)).get("files") == manifest and all(
                    os.path.isfile(os.path.join(DEST, r)) for r in FILES):
                return DEST
        except Exception:
            pass
    for rel in FILES:
        dst = os.path.join(DEST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(SRC, rel), dst)
    with open(mpath, "w") as f:
        json.dump({"source": SRC, "files": manifest}, f, indent=1)
    return DEST


if __name__ == "__main__":
    print(vendor(force="--force" in sys.argv))
    for rel in FILES:
        print(" ", rel)
