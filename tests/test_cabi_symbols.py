"""CPU test: the C-ABI library loads without a GPU and exports every symbol include/kbmod_b200.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "kbmod_b200.h")).read()
    return sorted(set(re.findall(r"KB_EXPORT\s+(?:const\s+char\s*\*|int)\s*(kb_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from kbmod_b200 import _lib
    syms = header_symbols()
    assert len(syms) >= 30
    dll = ctypes.CDLL(_lib.LIB_PATH)
    for s in syms:
        assert hasattr(dll, s), f"{s} declared in the header but not exported"
    assert sorted(_lib.SYMBOLS) == syms, "kbmod_b200/_lib.py SYMBOLS out of sync with the header"


def test_no_gpu_is_reported_not_faked():
    from kbmod_b200 import _lib
    n = _lib.device_count()
    assert n >= 0
    assert _lib.lib.kb_version().decode().startswith("kbmod_b200")


def test_struct_sizes_match_reference_records():
    from kbmod_b200 import _lib
    assert ctypes.sizeof(_lib.KbTrajectory) == 28  # tests/test_trajectory_list.py:28,47 in the reference
    assert _lib.TRAJECTORY_DTYPE.itemsize == 28
