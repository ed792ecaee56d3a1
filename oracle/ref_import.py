"""TEST INFRASTRUCTURE ONLY (oracle side) -- import the UNMODIFIED reference package from /root/reference.

The reference (t0m1ab/alphazero) is pure Python; it imports here with three shims:
  * `aenum` is not installed     -> a stub module re-exporting stdlib `enum.Enum` (+ a dummy NoAlias);
  * `matplotlib` is not installed -> empty stub modules (only the display code touches them);
  * `alphazero/__init__.py:1-2` asks importlib.metadata for the installed version -> patched to "0.1.0".
Nothing under /root/reference is modified or copied.  This module is used ONLY by the fixture generator
(tests/golden/make_golden.py) and by optional live-differential tests that skip when /root/reference is absent
(it does not exist on the GPU box).
"""
import importlib.metadata
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("AZ_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "alphazero"))


def import_reference():
    """Returns the imported reference top-level package `alphazero` (raises ImportError if absent)."""
    if not reference_available():
        raise ImportError(f"reference not present at {REFERENCE_ROOT}")
    if "aenum" not in sys.modules:
        try:
            import aenum  # noqa: F401
        except ImportError:
            import enum
            m = types.ModuleType("aenum")
            m.Enum = enum.Enum
            m.NoAlias = object()
            sys.modules["aenum"] = m
    if "matplotlib" not in sys.modules:
        try:
            import matplotlib  # noqa: F401
        except ImportError:
            mpl = types.ModuleType("matplotlib")
            mpl.__path__ = []
            for sub in ("pyplot", "patches", "colors"):
                sm = types.ModuleType(f"matplotlib.{sub}")
                setattr(mpl, sub, sm)
                sys.modules[f"matplotlib.{sub}"] = sm
            sys.modules["matplotlib.colors"].CSS4_COLORS = {}
            sys.modules["matplotlib.colors"].to_rgb = lambda c: (0.0, 0.0, 0.0)
            sys.modules["matplotlib"] = mpl
    _orig_version = importlib.metadata.version

    def _version(name):
        if name == "alphazero":
            return "0.1.0"
        return _orig_version(name)

    importlib.metadata.version = _version
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import alphazero  # the reference package
    return alphazero
