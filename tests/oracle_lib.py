"""TEST INFRASTRUCTURE: ctypes binding of the CPU oracle (oracle/libbrb_port.so) and of the compiled
reference binary (oracle/_ref/bramble_ref).  Never imported by the product package."""
import ctypes as C
import os
import subprocess

import numpy as np

from bramble_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PORT = os.path.join(ROOT, "oracle", "libbrb_port.so")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "bramble_ref")
KSW_REF = os.path.join(ROOT, "oracle", "_ref", "libksw_ref.so")      # the reference's own ksw2 translation unit (oracle/Makefile: kswref)


class PortOut(C.Structure):
    _fields_ = [("n_rec", C.c_uint32), ("n_cigar_words", C.c_uint64), ("group_rec_off", capi.u32p),
                ("src_aln", capi.u32p), ("tid", capi.u32p), ("pos", capi.u32p), ("tx_strand", capi.i8p),
                ("flags", capi.u8p), ("sim_score", capi.f64p), ("clip_score", capi.i32p), ("cigar_off", capi.u32p),
                ("cigar_len", capi.u32p), ("cigar", capi.u32p), ("n_matches", C.c_uint64), ("n_rescue", C.c_uint64),
                ("n_hits", C.c_uint64), ("seconds_project", C.c_double)]


_port = None


def port():
    global _port
    if _port is None:
        lib = C.CDLL(PORT)
        lib.brb_port_project.argtypes = [C.POINTER(capi.IndexDesc), C.POINTER(capi.Params), C.POINTER(capi.BatchIn), C.POINTER(PortOut)]
        lib.brb_port_free.argtypes = [C.POINTER(PortOut)]
        lib.brb_port_lookup.argtypes = [C.POINTER(capi.IndexDesc), C.POINTER(capi.Params), C.c_uint32, capi.i32p, capi.i8p,
                                        capi.u32p, capi.u32p, capi.u8p, C.c_uint32, capi.u32p, capi.u32p]
        lib.brb_port_merge_cigars.argtypes = [C.c_uint32, capi.u32p, capi.u32p, capi.u32p, capi.u32p, capi.u32p, capi.u32p, capi.u32p]
        lib.brb_port_ksw_extend.argtypes = [C.c_uint32, capi.u64p, capi.u8p, capi.u64p, capi.u8p, capi.i32p, capi.i32p, capi.i32p,
                                            capi.i32p, capi.u8p, C.c_uint32, capi.u32p, capi.u32p]
        lib.brb_port_flush_tags.argtypes = [C.c_uint32, capi.u32p, capi.u8p, capi.f64p, capi.u64p, capi.u8p, C.c_int, capi.i32p, capi.i32p, capi.u8p, capi.u8p]
        _port = lib
    return _port


def port_flush_tags(rows, batch, long_reads):
    """HI / NH / MAPQ / primary of the rows of `port_project` (emit order) with the real libstdc++ hash and RNG."""
    n = int(rows["n_rec"])
    hi = np.zeros(max(n, 1), np.int32); nh = np.zeros(max(n, 1), np.int32); mq = np.zeros(max(n, 1), np.uint8); pr = np.zeros(max(n, 1), np.uint8)
    gro = np.ascontiguousarray(rows["group_rec_off"], dtype=np.uint32)
    fl = np.ascontiguousarray(rows["flags"] if n else np.zeros(1, np.uint8), dtype=np.uint8)
    sim = np.ascontiguousarray(rows["sim_score"] if n else np.zeros(1), dtype=np.float64)
    rc = port().brb_port_flush_tags(int(batch.desc.n_groups), capi.ptr(gro), capi.ptr(fl), capi.ptr(sim), capi.ptr(batch.arrays["group_name_off"]),
                                    capi.ptr(batch.arrays["name_bytes"]), int(long_reads), capi.ptr(hi), capi.ptr(nh), capi.ptr(mq), capi.ptr(pr))
    assert rc == 0
    return dict(hi=hi[:n], nh=nh[:n], mapq=mq[:n], primary=pr[:n])


def rows_from(o, n_groups):
    """Copy a brb_batch_out / brb_port_out into plain numpy arrays (dict)."""
    n = int(o.n_rec)

    def arr(p, cnt, dt):
        if cnt == 0:
            return np.zeros(0, dtype=dt)
        return np.ctypeslib.as_array(p, shape=(cnt,)).astype(dt, copy=True)
    d = dict(n_rec=n,
             group_rec_off=arr(o.group_rec_off, n_groups + 1, np.uint32), src_aln=arr(o.src_aln, n, np.uint32),
             tid=arr(o.tid, n, np.uint32), pos=arr(o.pos, n, np.uint32), tx_strand=arr(o.tx_strand, n, np.int8),
             flags=arr(o.flags, n, np.uint8), sim_score=arr(o.sim_score, n, np.float64),
             clip_score=arr(o.clip_score, n, np.int32), cigar_off=arr(o.cigar_off, n, np.uint32),
             cigar_len=arr(o.cigar_len, n, np.uint32))
    pool = arr(o.cigar, int(o.n_cigar_words), np.uint32)
    d["cigars"] = [pool[int(a):int(a) + int(l)].copy() for a, l in zip(d["cigar_off"], d["cigar_len"])]
    return d


def port_project(index, params, batch):
    o = PortOut()
    rc = port().brb_port_project(C.byref(index.desc), C.byref(params), C.byref(batch.desc), C.byref(o))
    assert rc == 0
    d = rows_from(o, batch.desc.n_groups)
    d["n_matches"], d["n_rescue"], d["n_hits"] = int(o.n_matches), int(o.n_rescue), int(o.n_hits)
    port().brb_port_free(C.byref(o))
    return d


def run_reference(workdir, in_bam, gtf, out_bam, args=(), env=None, threads=1):
    """Runs the compiled, unmodified reference (it writes ./tmp_header.sam into its cwd)."""
    e = dict(os.environ)
    e.update(env or {})
    cmd = [REF_BIN, in_bam, "-G", gtf, "-o", out_bam, "-p", str(threads), "--quiet"] + list(args)
    r = subprocess.run(cmd, cwd=workdir, env=e, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    return r


def assert_rows_equal(a, b, what=""):
    assert a["n_rec"] == b["n_rec"], f"{what}: n_rec {a['n_rec']} != {b['n_rec']}"
    for k in ("group_rec_off", "src_aln", "tid", "pos", "tx_strand", "flags", "clip_score"):
        if not np.array_equal(a[k], b[k]):
            bad = np.nonzero(a[k] != b[k])[0][:5]
            raise AssertionError(f"{what}: column {k} differs at rows {bad}: {a[k][bad]} vs {b[k][bad]}")
    # similarity: bit-exact is the goal; north_star tolerance is 1e-6 relative
    if not np.array_equal(a["sim_score"], b["sim_score"]):
        np.testing.assert_allclose(a["sim_score"], b["sim_score"], rtol=1e-6, atol=0, err_msg=what)
    for r, (x, y) in enumerate(zip(a["cigars"], b["cigars"])):
        assert np.array_equal(x, y), f"{what}: cigar of row {r} differs: {x} vs {y}"
