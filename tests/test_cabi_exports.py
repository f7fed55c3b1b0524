"""CPU-only: the C-ABI library builds, loads, and exports every symbol include/lmi_b200.h declares.
No compute calls (there is no GPU here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import lmi_b200.build as b
    b.build()
    from lmi_b200 import _lib
    return _lib.load()


def test_header_symbols_are_exported(lib):
    header = open(os.path.join(ROOT, "include", "lmi_b200.h")).read()
    declared = set(re.findall(r"\b(lmi_[a-z_0-9]+)\s*\(", header))
    from lmi_b200._lib import SYMBOLS
    assert declared == set(SYMBOLS), declared ^ set(SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), f"{name} not exported"


def test_version_and_error_string(lib):
    assert b"sm_100a" in lib.lmi_version()
    assert isinstance(lib.lmi_last_error(), bytes)


def test_stats_struct_matches_header():
    from lmi_b200._lib import LmiStats
    # 4*int32 + 6*int64 + double + 6*float + int32 (+4 pad) + int64 + 2*float = 16 + 48 + 8 + 24 + 4 + 4 + 8 + 8
    assert ctypes.sizeof(LmiStats) == 120


def test_sass_is_blackwell_native():
    """The shipped .so must contain tcgen05 / TMA SASS (UTCHMMA, UTMALDG, LDTM), not legacy HMMA."""
    import shutil
    import subprocess
    from lmi_b200._lib import LIB_PATH
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", LIB_PATH], capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass and "UTMALDG" in sass and "LDTM" in sass
    assert "sm_100a" in sass
    assert not re.search(r"\bHMMA\b", sass)


def test_no_oracle_import_in_product():
    """The product package must never import the oracle (parity claims depend on it)."""
    pkg = os.path.join(ROOT, "sisap23-laion-challenge-learned-index_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(root, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f
