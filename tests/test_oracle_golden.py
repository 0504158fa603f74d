"""CPU: the oracle (oracle/libbrb_port.so) against every golden vector the reference's own tests hold
for this path."""
import ctypes as C
import re

import os

import numpy as np
import pytest

from bramble_b200 import capi, pack, synth
import oracle_lib

_OP = {c: i for i, c in enumerate("MIDNSHP=X")}
_OP.update({"m": 10, "d": 11, "i": 12, "s": 13})   # lower case = override codes 10..13 (include/evaluate.h:10-13)


def cig(s):
    return np.array([(int(n) << 4) | _OP[o] for n, o in re.findall(r"(\d+)([A-Za-z=])", s)], dtype=np.uint32)


# bramble-rs/tests/update_cigar.rs:21-143 (real, ideal, merged); valid for C++ merge_cigars
UPDATE_CIGAR_CASES = [
    ("2S8M", "2m8M", "10M"),
    ("10M", "5M1D4M", "5M1D5M"),
    ("10M", "5M1I5M", "5M1I5M"),
    ("2H10M", "10M", "2H10M"),
    ("10M", "5M3I3D5M", "5M3I3D5M"),
    ("5M3N5M", "10M", "10M"),
    ("10M", "2s8M", "2S8M"),
]


def port_merge(reals, ideals):
    n = len(reals)
    real_off = np.concatenate([[0], np.cumsum([len(x) for x in reals])]).astype(np.uint32)
    ideal_off = np.concatenate([[0], np.cumsum([len(x) for x in ideals])]).astype(np.uint32)
    real = np.concatenate(reals + [np.zeros(1, np.uint32)]).astype(np.uint32)
    ideal = np.concatenate(ideals + [np.zeros(1, np.uint32)]).astype(np.uint32)
    caps = np.diff(real_off.astype(np.int64)) + np.diff(ideal_off.astype(np.int64)) + 2
    out_off = np.concatenate([[0], np.cumsum(caps)]).astype(np.uint32)
    out_len = np.zeros(n, np.uint32); out = np.zeros(int(out_off[-1]) + 1, np.uint32)
    rc = oracle_lib.port().brb_port_merge_cigars(n, capi.ptr(real_off), capi.ptr(real), capi.ptr(ideal_off), capi.ptr(ideal),
                                                 capi.ptr(out_off), capi.ptr(out_len), capi.ptr(out))
    assert rc == 0
    return [out[int(o):int(o) + int(l)].copy() for o, l in zip(out_off[:-1], out_len)]


def test_update_cigar_golden():
    got = port_merge([cig(r) for r, _, _ in UPDATE_CIGAR_CASES], [cig(i) for _, i, _ in UPDATE_CIGAR_CASES])
    for g, (_, _, want) in zip(got, UPDATE_CIGAR_CASES):
        assert np.array_equal(g, cig(want))


def test_presets_golden():
    # bramble-rs/tests/evaluate.rs:6-56 == src/evaluate.cpp:1144-1156,1184-1189
    s = capi.resolve_params()
    assert (s.max_clip, s.max_junc_ins, s.max_junc_gap, s.max_error_exon) == (5, 0, 0, 0)
    assert s.similarity_threshold == 1.0 and not s.filter_by_similarity and not s.ignore_small_exons
    assert capi.resolve_params(strict=True).max_clip == 0
    l = capi.resolve_params(lr=True)
    assert (l.max_clip, l.max_junc_ins, l.max_junc_gap, l.max_error_exon) == (40, 40, 40, 35)
    assert abs(l.similarity_threshold - 0.60) < 1e-7 and l.filter_by_similarity and l.ignore_small_exons and l.long_reads
    h = capi.resolve_params(lr_hq=True, strict=True)       # STRICT branch unreachable for long reads
    assert (h.max_clip, h.max_junc_ins, h.max_junc_gap, h.max_error_exon) == (5, 10, 10, 35)
    assert abs(h.similarity_threshold - 0.90) < 1e-7
    assert capi.resolve_params(lr=True, lr_hq=True).max_clip == 40
    assert capi.resolve_params(max_clip=7, similarity_threshold=0.5).filter_by_similarity


def fixture_annotation():
    # bramble-rs/tests/short_read_projection.rs:41-56
    tx = [dict(id="tx1", gene="g1", chrom=0, strand="+", exons=[(101, 300)], family=0),
          dict(id="tx2", gene="g2", chrom=0, strand="+", exons=[(500, 600), (800, 900)], family=1)]
    return synth.Annotation(["chr1"], [10000], tx)


def fixture_reads():
    def rec(name, pos1, cg):
        return dict(name=name, flag=0, chrom=0, pos=pos1 - 1, mapq=60, cigar=cg, mchrom=-1, mpos=-1, isize=0, seq=b"A" * 100,
                    qual=b"I" * 100, tags=[])
    return [rec("r1", 151, [("M", 100)]), rec("r2", 551, [("M", 50), ("N", 199), ("M", 50)])]


def test_short_read_projection_fixture():
    # exact C++ answers derived in SURVEY.md 8(c).3
    ann = fixture_annotation()
    ix = pack.build_index(ann)
    assert ix.tid_names == ["tx1", "tx2"] and ix.tid_len == [200, 202]
    b = pack.pack_batches(fixture_reads(), ix, ann.chroms)[0]
    rows = oracle_lib.port_project(ix, capi.resolve_params(), b)
    assert rows["n_rec"] == 2
    assert list(rows["tid"]) == [0, 1] and list(rows["pos"]) == [50, 51]
    assert [list(c) for c in rows["cigars"]] == [[100 << 4], [100 << 4]]
    assert list(rows["flags"]) == [capi.REC_FIRST, capi.REC_FIRST]
    assert list(rows["group_rec_off"]) == [0, 1, 2]


KSW_T = "ACGTACGTTTGACCATGACCATGGGACGATCAGCATCGACTAGCATCGACTACGACTCGATCGACTAGCTAGCTAG"
KSW_Q = "ACGTACGTTGACCATGACCTTGGGACGATCAGCAAATCGACTAGCATCGACTACG"


def codes(s):
    return np.array(["ACGT".find(c) if c in "ACGT" else 4 for c in s], dtype=np.uint8)


def port_ksw(queries, targets, cap=256, fn=None):
    n = len(queries)
    q_off = np.concatenate([[0], np.cumsum([len(q) for q in queries])]).astype(np.uint64)
    t_off = np.concatenate([[0], np.cumsum([len(t) for t in targets])]).astype(np.uint64)
    q = np.ascontiguousarray(np.concatenate(list(queries) + [np.zeros(1, np.uint8)]).astype(np.uint8))
    t = np.ascontiguousarray(np.concatenate(list(targets) + [np.zeros(1, np.uint8)]).astype(np.uint8))
    mx = np.zeros(n, np.int32); sc = np.zeros(n, np.int32); mq = np.zeros(n, np.int32); mt = np.zeros(n, np.int32)
    zd = np.zeros(n, np.uint8); nc = np.zeros(n, np.uint32); cg = np.zeros((n, cap), np.uint32)
    rc = (fn or oracle_lib.port().brb_port_ksw_extend)(n, capi.ptr(q_off), capi.ptr(q), capi.ptr(t_off), capi.ptr(t), capi.ptr(mx), capi.ptr(sc),
                                               capi.ptr(mq), capi.ptr(mt), capi.ptr(zd), cap, capi.ptr(nc), capi.ptr(cg.reshape(-1)))
    assert rc == 0
    return dict(max=mx, score=sc, max_q=mq, max_t=mt, zdropped=zd, n_cigar=nc, cigars=[cg[i, :min(int(nc[i]), cap)].copy() for i in range(n)])


def test_ksw2_smoke_vector():
    # reference TU ksw2_extz2_sse.cpp with bramble's parameters
    r = port_ksw([codes(KSW_Q)], [codes(KSW_T)])
    assert (int(r["max"][0]), int(r["score"][0]), int(r["max_t"][0]), int(r["max_q"][0]), int(r["zdropped"][0])) == (37, 11, 53, 54, 0)
    ops = "".join(f"{int(w) >> 4}{'MID'[int(w) & 15]}" for w in r["cigars"][0])
    assert ops == "7M1D26M2I20M"


@pytest.mark.skipif(not os.path.exists(oracle_lib.KSW_REF), reason="oracle/_ref/libksw_ref.so not built (no reference sources)")
def test_ksw_port_matches_reference_translation_unit():
    """The restatement of ksw_extz2_sse + ksw_backtrack (oracle/brb_port.cpp, oracle/shim/ksw2.h) against the reference's OWN
    translation unit (subprojects/packagefiles/ksw2/ksw2_extz2_sse.cpp, compiled unmodified) on the problems the GPU kernel
    is tested on plus 2000 more: max, score (incl. NEG_INF), end cell, z-drop flag and CIGAR, problem by problem."""
    import ctypes as C
    import common
    lib = C.CDLL(oracle_lib.KSW_REF)
    lib.brb_ref_ksw_extend.argtypes = oracle_lib.port().brb_port_ksw_extend.argtypes
    qs, ts = [codes(KSW_Q)], [codes(KSW_T)]
    for seed, n, smax in ((17, 400, 120), (18, 1000, 60), (19, 1000, 400)):
        a, b = common.ksw_problems(seed, n, smax)
        qs += a; ts += b
    ref = port_ksw(qs, ts, cap=2048, fn=lib.brb_ref_ksw_extend)
    got = port_ksw(qs, ts, cap=2048)
    for k in ("max", "score", "max_q", "max_t", "zdropped", "n_cigar"):
        assert np.array_equal(got[k], ref[k]), k
    for i, (g, e) in enumerate(zip(got["cigars"], ref["cigars"])):
        assert np.array_equal(g, e), i
    assert (int(ref["max"][0]), int(ref["score"][0]), int(ref["max_t"][0]), int(ref["max_q"][0])) == (37, 11, 53, 54)
    assert ref["zdropped"].sum() > 100 and (ref["score"] > -0x40000000).sum() > 300 and (ref["score"] == -0x40000000).sum() > 300
