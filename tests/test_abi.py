"""The C-ABI library loads without a GPU and exports every symbol include/fetode.h declares."""
import ctypes
import os
import re

from conftest import PKG, ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "fetode.h")).read()
    return sorted(set(re.findall(r"FETODE_API\s+[\w\s\*]+?\b(fetode_\w+)\s*\(", src)))


def test_header_declares_entry_points():
    names = _declared()
    assert "fetode_smallnet_rk_fwd" in names and "fetode_kanlinear_fwd" in names
    assert len(names) >= 15


def test_library_exports_every_declared_symbol():
    import build as fetode_build                       # fet-ode_b200/build.py
    path = fetode_build.build()
    lib = ctypes.CDLL(path)
    missing = [n for n in _declared() if not hasattr(lib, n)]
    assert not missing, f"declared in include/fetode.h but not exported: {missing}"
    lib.fetode_version.restype = ctypes.c_int
    assert lib.fetode_version() >= 100
    lib.fetode_strerror.restype = ctypes.c_char_p
    assert b"shape" in lib.fetode_strerror(-4)
    # host-only queries of the allocation-free variant: its limit is the header's, and the error it answers with has a text
    src = open(os.path.join(ROOT, "include", "fetode.h")).read()
    cap = int(re.search(r"#define\s+FETODE_STATIC_WS_FLOATS\s+(\d+)", src).group(1))
    lib.fetode_kanlinear_bwd_static_max_floats.restype = ctypes.c_int64
    assert lib.fetode_kanlinear_bwd_static_max_floats() == cap == 262144
    toolarge = int(re.search(r"#define\s+FETODE_ETOOLARGE\s+\((-?\d+)\)", src).group(1))
    assert b"static" in lib.fetode_strerror(toolarge)


def test_every_reference_citation_in_header_names_a_file_and_line():
    src = open(os.path.join(ROOT, "include", "fetode.h")).read()
    cites = re.findall(r"(\w+\.py):(\d+)", src)
    assert len(cites) >= 20
