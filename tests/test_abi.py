"""CPU: the C-ABI library loads, exports every symbol include/bramble_b200.h declares, and refuses to
compute without a device (no CPU fallback)."""
import ctypes as C
import os
import re

from bramble_b200 import capi
import common

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "bramble_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(brb_[a-z0-9_]+)\s*\(", src)))


def test_all_declared_symbols_are_exported():
    lib = C.CDLL(capi.LIB_PATH)
    names = declared_functions()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/bramble_b200.h but not exported"


def test_struct_sizes_match_header():
    assert C.sizeof(capi.Params) == 24
    assert C.sizeof(capi.BatchIn) == 8 + 11 * 8
    assert C.sizeof(capi.IndexDesc) == 8 + 12 * 8 + 8 + 2 * 8


def test_no_device_no_compute():
    lib = capi.lib()
    assert lib.brb_version().startswith(b"bramble_b200")
    if lib.brb_device_count() > 0:
        return
    wl = common.make_workload("default", 3, n_names=20, n_genes=10)
    ix = C.c_void_p()
    rc = lib.brb_index_create(0, C.byref(wl["index"].desc), C.byref(ix))
    assert rc == capi.BRB_E_NO_DEVICE
    assert b"device" in lib.brb_last_error()
    import pytest
    from bramble_b200 import engine
    with pytest.raises(capi.EngineError):
        engine.Engine(wl["index"])
