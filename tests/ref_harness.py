"""
Loads the UNMODIFIED reference sources from /root/reference with the oracle/pyg_shim stand-ins
for its missing third-party imports (builder container only; the GPU box has no /root/reference).
"""
import contextlib
import io
import os
import sys
import types

REF_SRC = "/root/reference/src/nichecompass"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "oracle", "pyg_shim")


def reference_available() -> bool:
    return os.path.isdir(REF_SRC)


def load_reference():
    """Returns the reference ``nichecompass`` package object (sub-packages nn, modules, data,
    train imported; ``nichecompass/__init__.py`` itself is skipped because it pulls plotting /
    benchmarking extras and importlib.metadata.version)."""
    if "nichecompass" in sys.modules and getattr(sys.modules["nichecompass"], "_is_ref", False):
        return sys.modules["nichecompass"]
    for p in (ROOT, SHIM):
        if p not in sys.path:
            sys.path.insert(0, p)
    pkg = types.ModuleType("nichecompass")
    pkg.__path__ = [REF_SRC]
    pkg._is_ref = True
    sys.modules["nichecompass"] = pkg
    import importlib
    for sub in ("nn", "modules"):
        setattr(pkg, sub, importlib.import_module(f"nichecompass.{sub}"))
    return pkg


def load_reference_full():
    """Additionally imports nichecompass.data / nichecompass.train (needs the loader shims)."""
    pkg = load_reference()
    import importlib
    # data/__init__ imports datareaders (pandas/anndata only) -- fine with the stand-ins
    for sub in ("data", "train"):
        setattr(pkg, sub, importlib.import_module(f"nichecompass.{sub}"))
    return pkg


@contextlib.contextmanager
def quiet():
    with contextlib.redirect_stdout(io.StringIO()):
        yield
