"""TEST INFRASTRUCTURE: loader for the reference build in ``oracle/_ref`` (see ``build_ref.sh``).

Only ``tests/``, ``tests/golden/make_golden.py``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this module.  The product package ``kbmod_b200``
never does.
"""
import glob
import importlib.machinery
import importlib.util
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")


def _so_path():
    hits = sorted(glob.glob(os.path.join(_REF_DIR, "search*.so")))
    return hits[0] if hits else None


def available():
    return _so_path() is not None


def _load(name):
    if name in sys.modules:
        return sys.modules[name]
    path = _so_path()
    if path is None:
        raise ImportError("oracle/_ref is not built; run oracle/build_ref.sh where /root/reference exists")
    loader = importlib.machinery.ExtensionFileLoader(name, path)
    spec = importlib.util.spec_from_file_location(name, path, loader=loader)
    mod = importlib.util.module_from_spec(spec)
    loader.exec_module(mod)
    sys.modules[name] = mod
    return mod


def load():
    """Return (search, ref_harness): the reference's own ``kbmod.search`` module and our harness."""
    search = _load("search")
    harness = _load("ref_harness")
    return search, harness
