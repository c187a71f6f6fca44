"""The reference's OWN acceptance tests for the driver-side functions of the path, run UNMODIFIED against verl_b200.

The files are the snapshot made by ``python -m oracle.vendor_ref`` (oracle/_ref/tests/..., git-ignored, travels to the
GPU box): tests/trainer/ppo/test_core_algos_on_cpu.py (registry semantics + multi-turn GAE invariance on CPU tensors)
and tests/utils/test_torch_functional.py::test_masked_mean (CPU tensors, NaN outside the mask).  They import
``verl.trainer.ppo.core_algos`` / ``verl.utils.torch_functional``; the ``vendored_install`` plugin points those paths at
verl_b200 (compat.install()).  The tests pass HOST tensors exactly like the reference's driver does
(verl/trainer/ppo/ray_trainer.py:1283-1305): verl_b200.staging moves them to the device and back."""

import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_TESTS = os.path.join(ROOT, "oracle", "_ref", "tests")

needs_snapshot = pytest.mark.skipif(not os.path.isfile(os.path.join(ROOT, "oracle", "_ref", "MANIFEST.json")),
                                    reason="oracle/_ref snapshot missing (python -m oracle.vendor_ref)")


def _run(args, plugin="vendored_install"):
    env = dict(os.environ, PYTHONPATH=os.path.join(ROOT, "tests") + os.pathsep + ROOT + os.pathsep +
               os.environ.get("PYTHONPATH", ""))
    cmd = [sys.executable, "-m", "pytest", "-q", "-p", plugin, "-p", "no:cacheprovider",
           "--rootdir", REF_TESTS, "-c", os.devnull, *args]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env, cwd=REF_TESTS)


@needs_snapshot
def test_reference_registry_tests_run_on_verl_b200():
    """TestRegisterAdvEst (9 tests): pure host logic of register_adv_est / get_adv_estimator_fn — no device needed."""
    r = _run([os.path.join(REF_TESTS, "trainer", "ppo", "test_core_algos_on_cpu.py"), "-k", "TestRegisterAdvEst"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "9 passed" in r.stdout, r.stdout[-2000:]


@needs_snapshot
@pytest.mark.gpu
def test_reference_cpu_acceptance_tests_on_device():
    """The whole of test_core_algos_on_cpu.py (10 tests) + test_masked_mean (3 cases), host tensors in and out."""
    r = _run([os.path.join(REF_TESTS, "trainer", "ppo", "test_core_algos_on_cpu.py"),
              os.path.join(REF_TESTS, "utils", "test_torch_functional.py"), "-k", "not distributed"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "13 passed" in r.stdout, r.stdout[-2000:]


@needs_snapshot
@pytest.mark.gpu
def test_reference_acceptance_tests_with_importable_reference_patched():
    """Deployment mode: the reference's own module objects are importable (loaded from the snapshot) and
    compat.install() PATCHES the hot-path names into them.  The reference's core_algos tests (10), masked_mean cases (3)
    and its compute_data_metrics tests (tests/trainer/ppo/test_metric_utils_on_cpu.py::TestComputeDataMetrics, 2 — a
    MagicMock batch of host tensors) run unmodified on top."""
    r = _run([os.path.join(REF_TESTS, "trainer", "ppo", "test_core_algos_on_cpu.py"),
              os.path.join(REF_TESTS, "utils", "test_torch_functional.py"),
              os.path.join(REF_TESTS, "trainer", "ppo", "test_metric_utils_on_cpu.py"),
              "-k", "not distributed and (TestComputeDataMetrics or not metric_utils)"], plugin="vendored_install_patched")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "15 passed" in r.stdout, r.stdout[-2000:]


@needs_snapshot
def test_patched_mode_plugin_patches_reference_modules():
    """No device needed: with the reference's modules importable, install() reports 'patched' for them and replaces the
    names (the plugin asserts it at import); the registry tests of the reference still pass on the patched module."""
    r = _run([os.path.join(REF_TESTS, "trainer", "ppo", "test_core_algos_on_cpu.py"), "-k", "TestRegisterAdvEst"],
             plugin="vendored_install_patched")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "9 passed" in r.stdout, r.stdout[-2000:]
