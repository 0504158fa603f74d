"""brb_merge_cigars_static -- the CIGAR merge with its scratch in a fixed __device__ array of the library (no cudaMalloc in the call,
no counterpart in the reference) -- against the existing allocating entry point
brb_merge_cigars, through the C ABI, on two calls in a row that share the array.  Bit-exact: CIGAR words are integers.

Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from bramble_b200 import capi, engine
import common
from test_gpu_parity import _merge_on_gpu, _random_cigar_pairs

pytestmark = pytest.mark.gpu


def _dev(a):
    """numpy uint32 array -> device tensor (int32 view: the bits are what matters)"""
    import torch
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint32).view(np.int32).copy()).cuda()


def _static_merge(lib, reals, ideals, max_real, max_ideal, stream=None):
    import torch
    n = len(reals)
    real_off = np.concatenate([[0], np.cumsum([len(x) for x in reals])]).astype(np.uint32)
    ideal_off = np.concatenate([[0], np.cumsum([len(x) for x in ideals])]).astype(np.uint32)
    caps = np.array([len(r) + len(i) + 2 for r, i in zip(reals, ideals)], np.int64)
    out_off = np.concatenate([[0], np.cumsum(caps)]).astype(np.uint32)
    d_ro, d_io, d_oo = _dev(real_off), _dev(ideal_off), _dev(out_off)
    d_r = _dev(np.concatenate(reals + [np.zeros(1, np.uint32)]))
    d_i = _dev(np.concatenate(ideals + [np.zeros(1, np.uint32)]))
    d_len = torch.full((max(n, 1),), -1, dtype=torch.int32, device="cuda")
    d_out = torch.full((int(out_off[-1]) + 1,), -1, dtype=torch.int32, device="cuda")
    d_ns = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc = lib.brb_merge_cigars_static(0, stream, n, max_real, max_ideal, d_ro.data_ptr(), d_r.data_ptr(), d_io.data_ptr(), d_i.data_ptr(),
                                     d_oo.data_ptr(), d_len.data_ptr(), d_out.data_ptr(), d_ns.data_ptr())
    if rc != 0:
        return rc, None, None
    torch.cuda.synchronize()
    lens = d_len.cpu().numpy().view(np.uint32)
    out = d_out.cpu().numpy().view(np.uint32)
    return 0, [out[int(o):int(o) + int(l)].copy() for o, l in zip(out_off[:-1], lens[:n])], int(d_ns.cpu().numpy()[0])


def test_static_scratch_merge_matches_allocating_entry_point():
    import torch
    lib = capi.lib()
    wl = common.make_workload("default", 83, n_names=20)
    # call 1: long CIGARs (up to 700 real / 60 ideal ops: slices of ~1000 words), incl. zero-length ops (the serial hand-over inside the warp code)
    r1, i1 = _random_cigar_pairs(21, 300, 700, 60)
    ra, ia = _random_cigar_pairs(22, 300, 120, 20, True)
    r1 += ra; i1 += ia
    # call 2, right behind it on another stream: short CIGARs with bounds that are too small for some pairs (they take the serial path and are counted)
    r2, i2 = _random_cigar_pairs(23, 4000, 60, 12)
    with engine.Engine(wl["index"]) as eng:
        want1 = _merge_on_gpu(eng, r1, i1, True)
        want2 = _merge_on_gpu(eng, r2, i2, False)
        rc, got1, ns1 = _static_merge(lib, r1, i1, 700, 60)
        assert rc == 0, lib.brb_last_error()
        st = torch.cuda.Stream()
        rc, got2, ns2 = _static_merge(lib, r2, i2, 40, 8, stream=st.cuda_stream)
        assert rc == 0, lib.brb_last_error()
        # ... and the first data again, after the array has held the second call's contents
        rc, got3, ns3 = _static_merge(lib, r1, i1, 700, 60)
        assert rc == 0, lib.brb_last_error()
    for k, (g, e) in enumerate(zip(got1, want1)):
        assert np.array_equal(g, e), ("call 1", k)
    for k, (g, e) in enumerate(zip(got2, want2)):
        assert np.array_equal(g, e), ("call 2", k)
    for k, (g, e) in enumerate(zip(got3, want1)):
        assert np.array_equal(g, e), ("call 3", k)
    too_big = sum(1 for r, i in zip(r2, i2) if len(r) + 4 * len(i) + 16 > 40 + 4 * 8 + 16)
    assert ns1 == 0 and ns3 == 0 and ns2 == too_big and too_big > 100
    # sizes beyond the array are refused with a clear message; nothing is launched
    rc, _, _ = _static_merge(lib, r2[:4], i2[:4], 1 << 20, 8)
    assert rc == capi.E_ARG if hasattr(capi, "E_ARG") else rc != 0
    assert b"exceed the static scratch" in lib.brb_last_error()
    rc, got0, _ = _static_merge(lib, [], [], 10, 10)
    assert rc == 0 and got0 == []
