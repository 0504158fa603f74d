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
    assert C.sizeof(capi.BatchIn) == 8 + 13 * 8 + 5 * 8
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


def test_ctypes_mirrors_match_the_compiled_header(tmp_path):
    """The ctypes structures are hand-written mirrors: compile the header with gcc and compare sizes and the offsets
    of the last fields (a field added to one side only moves them)."""
    import subprocess
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "bramble_b200.h"\n'
                   'int main(void) { printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(brb_params), sizeof(brb_batch_in), sizeof(brb_index_desc),\n'
                   '  sizeof(brb_batch_out), sizeof(brb_stats), offsetof(brb_batch_out, packed_rows), offsetof(brb_batch_out, aln_rec_off),\n'
                   '  offsetof(brb_stats, alg_bytes_front)); return 0; }\n')
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    want = [C.sizeof(capi.Params), C.sizeof(capi.BatchIn), C.sizeof(capi.IndexDesc), C.sizeof(capi.BatchOut), C.sizeof(capi.Stats),
            capi.BatchOut.packed_rows.offset, capi.BatchOut.aln_rec_off.offset, capi.Stats.alg_bytes_front.offset]
    assert got == want
    assert (capi.OPT_GENERIC_FRONT, capi.OPT_NO_FASTPATH, capi.OPT_ORDERED_ROWS, capi.OPT_PACKED_ROWS) == (1, 2, 3, 4)
