"""pytest plugin (``-p vendored_install``) for the reference's own test files under oracle/_ref/tests: routes the
reference's import paths to verl_b200 BEFORE the test modules are imported."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import verl_b200.compat as _compat  # noqa: E402

REPORT = _compat.install()
