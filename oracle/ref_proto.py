"""Imports the reference's numpy prototypes IN PLACE from /root/reference (this container only).

Used by tests/golden/make_golden.py to pin the oracle; never used at test/bench run time on the
GPU box (where /root/reference does not exist).  matplotlib is absent, so a stub module is
registered before import; the prototypes' drivers sit under `if __name__ == "__main__"`.
"""
from __future__ import annotations

import importlib.util
import io
import contextlib
import os
import sys
import types

REF_ROOT = os.environ.get("TLFEA_REFERENCE_ROOT", "/root/reference")
T10_DIR = os.path.join(REF_ROOT, "test-scripts", "T10-tets")


def available() -> bool:
    return os.path.isdir(T10_DIR)


def _stub_matplotlib():
    if "matplotlib" not in sys.modules:
        m = types.ModuleType("matplotlib")
        p = types.ModuleType("matplotlib.pyplot")
        m.pyplot = p
        sys.modules["matplotlib"] = m
        sys.modules["matplotlib.pyplot"] = p


def load(script: str):
    """script: file name under test-scripts/T10-tets, e.g. 'f-form-T10-beam-newton.py'."""
    _stub_matplotlib()
    if T10_DIR not in sys.path:
        sys.path.insert(0, T10_DIR)
    path = os.path.join(T10_DIR, script)
    name = "ref_" + script.replace("-", "_").replace(".py", "")
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    with contextlib.redirect_stdout(io.StringIO()):
        spec.loader.exec_module(mod)
    return mod


@contextlib.contextmanager
def quiet():
    with contextlib.redirect_stdout(io.StringIO()):
        yield
