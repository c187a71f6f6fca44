"""pytest plugin (``-p vendored_install_patched``): the deployment mode with an IMPORTABLE reference.  The reference's own
modules are loaded from the oracle/_ref snapshot first (oracle.ref_import: the real ``verl.trainer.ppo.core_algos``,
``verl.utils.torch_functional``, ``verl.trainer.ppo.metric_utils`` module objects, unmodified), then
``verl_b200.compat.install()`` patches the hot-path names into them attribute by attribute — everything else in those
modules stays the reference's.  The reference's test files then import ``verl.*`` as usual."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
os.environ["VERL_REFERENCE_ROOT"] = os.path.join(ROOT, "oracle", "_ref")  # the snapshot, also where /root/reference exists

from oracle import ref_import  # noqa: E402

ref_import.load()
sys.modules["verl"].DataProto = type("DataProto", (), {})  # type hint of metric_utils.py (`from verl import DataProto`)
import importlib  # noqa: E402

REF_METRIC_UTILS = importlib.import_module("verl.trainer.ppo.metric_utils")
REF_COMPUTE_DATA_METRICS = REF_METRIC_UTILS.compute_data_metrics

import verl_b200.compat as _compat  # noqa: E402

REPORT = _compat.install()
assert REPORT["verl.trainer.ppo.core_algos"] == "patched", REPORT
assert REPORT["verl.utils.torch_functional"] == "patched", REPORT
assert REPORT["verl.trainer.ppo.metric_utils"] == "patched", REPORT
assert REF_METRIC_UTILS.compute_data_metrics is not REF_COMPUTE_DATA_METRICS
