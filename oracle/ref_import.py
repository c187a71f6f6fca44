"""Load the reference's hot-path modules straight from /root/reference (build container only; TEST INFRASTRUCTURE).

`import verl` fails here (ray / tensordict / omegaconf / hydra are not installed), but the hot-path modules only need
`verl`, `verl.utils`, `verl.trainer` to exist as packages and two type-hint-only names
(`verl.trainer.config.AlgoConfig`, `tensordict.TensorDict`).  Those are registered as bare stubs; every function
body that runs is the reference's own, unmodified.
"""

from __future__ import annotations

import importlib
import os
import sys
import types

_SNAPSHOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")  # oracle/vendor_ref.py (travels to the GPU box)
REFERENCE_ROOT = os.environ.get("VERL_REFERENCE_ROOT") or (
    "/root/reference" if os.path.isdir("/root/reference/verl") else _SNAPSHOT)


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "verl", "trainer", "ppo"))


def _pkg(name: str, path: str) -> types.ModuleType:
    mod = sys.modules.get(name)
    if mod is None:
        mod = types.ModuleType(name)
        mod.__path__ = [path]
        sys.modules[name] = mod
        parent, _, leaf = name.rpartition(".")
        if parent:
            setattr(sys.modules[parent], leaf, mod)
    return mod


_loaded = None


def load():
    """-> namespace with .core_algos, .torch_functional, .fused_torch (experimental.torch_functional), .model"""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    # refuse to mix with verl_b200.compat shadows
    for name in list(sys.modules):
        if name == "verl" or name.startswith("verl."):
            mod = sys.modules[name]
            if getattr(mod, "__file__", None) and "verl_b200" in str(mod.__file__):
                del sys.modules[name]
            elif not getattr(mod, "__file__", None) and not getattr(mod, "__path__", None):
                del sys.modules[name]
    root = os.path.join(REFERENCE_ROOT, "verl")
    _pkg("verl", root)
    _pkg("verl.utils", os.path.join(root, "utils"))
    _pkg("verl.trainer", os.path.join(root, "trainer"))
    _pkg("verl.trainer.ppo", os.path.join(root, "trainer", "ppo"))
    _pkg("verl.utils.experimental", os.path.join(root, "utils", "experimental"))
    cfg = types.ModuleType("verl.trainer.config")
    cfg.AlgoConfig = type("AlgoConfig", (), {})
    sys.modules["verl.trainer.config"] = cfg
    sys.modules["verl.trainer"].config = cfg
    if "tensordict" not in sys.modules:
        td = types.ModuleType("tensordict")
        td.TensorDict = type("TensorDict", (), {})
        sys.modules["tensordict"] = td
    ns = types.SimpleNamespace()
    ns.torch_functional = importlib.import_module("verl.utils.torch_functional")
    ns.core_algos = importlib.import_module("verl.trainer.ppo.core_algos")
    ns.fused_torch = importlib.import_module("verl.utils.experimental.torch_functional")
    try:
        ns.model = importlib.import_module("verl.utils.model")
    except Exception:  # needs transformers internals; only create_random_mask is used
        ns.model = None
    _loaded = ns
    return ns


def load_kernel():
    """-> (kernels module, linear_cross_entropy module) of the reference's Triton op (needs triton; runs on a GPU only)."""
    load()
    _pkg("verl.utils.kernel", os.path.join(REFERENCE_ROOT, "verl", "utils", "kernel"))
    return (importlib.import_module("verl.utils.kernel.kernels"),
            importlib.import_module("verl.utils.kernel.linear_cross_entropy"))


def load_functions(rel_path: str, names, namespace: dict) -> dict:
    """Compile ONLY the named top-level function / class definitions of a reference file whose module cannot be imported here
    (ray / hydra / tensordict at import time) and execute them, unmodified, in ``namespace`` (which supplies the
    module-level names the bodies use).  -> {name: function}"""
    import ast

    path = os.path.join(REFERENCE_ROOT, rel_path)
    tree = ast.parse(open(path).read(), filename=path)
    picked = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in set(names)]
    missing = set(names) - {n.name for n in picked}
    if missing:
        raise RuntimeError(f"{rel_path}: functions not found: {sorted(missing)}")
    for top in picked:  # annotations may name types that are not importable here; the bodies never evaluate them
        for fn in ast.walk(top):
            if isinstance(fn, ast.FunctionDef):
                fn.returns = None
                for a in fn.args.args + fn.args.kwonlyargs:
                    a.annotation = None
    mod = ast.Module(body=picked, type_ignores=[])
    ns = dict(namespace)
    exec(compile(mod, path, "exec"), ns)
    return {n: ns[n] for n in names}
