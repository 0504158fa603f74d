#!/usr/bin/env python
"""Generates tests/golden/quirks/<name>/ -- one hand-built (GTF, BAM[, FASTA]) pair per deliberately preserved quirk of
the reference, with the expected output produced by the UNMODIFIED reference binary
(oracle/_ref/bramble_ref, built by oracle/Makefile from /root/reference).  Run here (the container that has the
reference); the fixtures are committed and the tests read only them:

    python tests/golden/make_quirks.py

Every fixture also asserts that the quirk is actually visible in the reference's output, so that a fixture cannot
silently stop exercising what it is named after.
"""
import os
import shutil
import struct
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from bramble_b200 import synth  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "bramble_ref")
OUT = os.path.join(HERE, "quirks")
TAGS = [("NH", "C", 1), ("HI", "C", 1), ("AS", "C", 100), ("NM", "C", 0)]


def rec(name, flag, pos1, cigar, mpos1=0, seq=None, tags=None, mapq=60, paired_chrom=True):
    """pos1 / mpos1: 1-based leftmost coordinates (0 = none)."""
    ql = sum(l for o, l in cigar if o in "MIS=X")
    if seq is None:
        seq = bytes(b"ACGT"[(i * 7 + len(name)) % 4] for i in range(ql))
    return dict(name=name, flag=flag, chrom=0, pos=pos1 - 1, mapq=mapq, cigar=cigar, mchrom=0 if (flag & 1 and paired_chrom) else -1,
                mpos=mpos1 - 1 if flag & 1 else -1, isize=0, seq=seq, qual=bytes([30]) * len(seq), tags=list(TAGS if tags is None else tags))


def tx(tid, strand, exons, gene=None):
    return dict(id=tid, gene=gene or tid.split(".")[0], chrom=0, strand=strand, exons=exons, family=0)


def parse(path):
    """[(name, flag, tid, pos0, cigar string, aux bytes)] of an output BAM."""
    _, refs, recs = synth.read_bam(path)
    out = []
    for r in recs:
        tid, pos, bmq, fnc, lq, mtid, mpos, isz = struct.unpack_from("<iiIIiiii", r, 0)
        ln, nc = bmq & 0xff, fnc & 0xffff
        name = r[32:32 + ln - 1].decode()
        cg = struct.unpack_from(f"<{nc}I", r, 32 + ln)
        cs = "".join(f"{c >> 4}{'MIDNSHP=X'[c & 15]}" for c in cg)
        aux = r[32 + ln + 4 * nc + (lq + 1) // 2 + lq:]
        out.append((name, fnc >> 16, refs[tid][0] if tid >= 0 else "*", pos, cs, aux))
    return out


def fixture(name, ann, recs, args, check, genome=None, doc=""):
    d = os.path.join(OUT, name)
    shutil.rmtree(d, ignore_errors=True)
    os.makedirs(os.path.join(d, "run"))
    synth.write_gtf(ann, os.path.join(d, "a.gtf"))
    synth.write_bam(os.path.join(d, "in.bam"), ann, recs)
    argv = ["bramble", "../in.bam", "-G", "../a.gtf", "-o", "out.bam", "-p", "1", "--quiet"] + list(args)
    if genome is not None:
        synth.write_fasta(ann, genome, os.path.join(d, "g.fa"))
        argv += ["-S", "../g.fa"]
    r = subprocess.run(argv, executable=REF, cwd=os.path.join(d, "run"), capture_output=True, text=True, env=dict(os.environ, BRB_SHIM_LEVEL="6"))
    assert r.returncode == 0, r.stderr
    os.replace(os.path.join(d, "run", "out.bam"), os.path.join(d, "expected.bam"))
    shutil.rmtree(os.path.join(d, "run"))
    with open(os.path.join(d, "args.txt"), "w") as f:
        f.write(" ".join(argv[9:]) + "\n")
    with open(os.path.join(d, "README"), "w") as f:
        f.write(doc.strip() + "\n")
    got = parse(os.path.join(d, "expected.bam"))
    check(got)
    print(f"{name}: {len(got)} output records")
    for g in got:
        print("   ", g[:5])


def main():
    assert os.path.exists(REF), "build the reference first: make -C oracle ref"
    os.makedirs(OUT, exist_ok=True)

    # 1 -- the mate is dropped when the FIRST-SEEN end has no match (src/core.cpp:387-393: nothing is emitted for the pair,
    #      although the second end matched); the other way round the matching end comes out unpaired (mates.cpp:153: mate == nullptr)
    ann = synth.Annotation(["chr1"], [300000], [tx("G1.t1", "+", [(1000, 1200), (2000, 2200)])])
    recs = [rec("p1", 99, 5000, [("M", 100)], 1050), rec("p1", 147, 1050, [("M", 100)], 5000),       # first end intergenic, second end in exon 1
            rec("p2", 99, 1060, [("M", 100)], 6000), rec("p2", 147, 6000, [("M", 100)], 1060)]       # first end in exon 1, second intergenic

    def chk(got):
        assert [g[0] for g in got] == ["p2"], got
        assert got[0][1] & 1 == 0 and got[0][2] == "G1.t1"
    fixture("mate_first_no_match", ann, recs, [], chk, doc="""
mate_first_no_match: pair p1 = (intergenic end first, exon end second) -> the reference emits NOTHING (src/core.cpp:387-393);
pair p2 = (exon end first, intergenic end second) -> the exon end is emitted as an unpaired record (src/mates.cpp:150-176).""")

    # 2 -- BAM_FMREVERSE is set from the record's OWN transcript strand (src/bam.cpp:549-555): on a '-' transcript both ends carry it
    ann = synth.Annotation(["chr1"], [300000], [tx("G2.t1", "-", [(10000, 10400), (11000, 11300)]), tx("G2.t2", "+", [(20000, 20400)])])
    recs = [rec("q1", 99, 10050, [("M", 100)], 10200), rec("q1", 147, 10200, [("M", 100)], 10050),
            rec("q2", 99, 20050, [("M", 100)], 20200), rec("q2", 147, 20200, [("M", 100)], 20050)]

    def chk(got):
        q1 = [g for g in got if g[0] == "q1"]
        q2 = [g for g in got if g[0] == "q2"]
        assert sorted(g[1] for g in q1) == [99 ^ 16, (147 ^ 16) | 32], q1    # '-' transcript: both ends flipped, MREVERSE on BOTH (the second end gains it)
        assert sorted(g[1] for g in q2) == [99, 147], q2                    # '+' transcript: flags as they came
    fixture("fmreverse_own_strand", ann, recs, [], chk, doc="""
fmreverse_own_strand: set_mate_info sets BAM_FMREVERSE iff the record's own transcript strand is '-' (src/bam.cpp:549-555), so both
ends of a pair on a '-' transcript carry it (the input's reverse end gains it although its mate now points forward); on a '+'
transcript the flag is left as it came.""")

    # 3 -- the group-shared read sequence is the FIRST record's that has one, whatever its orientation (src/core.cpp:353-378):
    #      record B (reverse flag, no sequence of its own) gets its clip rescued with record A's bases
    rng = np.random.default_rng(5)
    g = bytearray(bytes(b"ACGT"[x] for x in rng.integers(0, 4, size=40000)))
    E1, E2 = (20001, 20200), (21001, 21300)
    F1, F2 = (30001, 30200), (31001, 31300)
    seqA = bytes(g[E1[1] - 30:E1[1]]) + bytes(g[E2[0] - 1:E2[0] - 1 + 270])          # 30 bases of E1's end + 270 bases of E2
    g[F1[1] - 25:F1[1]] = seqA[0:25]                                                 # F1 ends with what A's sequence starts with
    ann = synth.Annotation(["chr1"], [40000], [tx("G3.t1", "+", [E1, E2]), tx("G4.t1", "+", [F1, F2])])
    lt = [("NM", "C", 0), ("AS", "i", 250), ("ts", "A", "+"), ("NH", "C", 2), ("HI", "C", 1)]
    recs = [rec("L1", 0, E2[0], [("S", 30), ("M", 270)], seq=seqA, tags=lt),
            rec("L1", 256 | 16, F2[0], [("S", 25), ("M", 200)], seq=b"", tags=lt)]

    def chk(got):
        a = [x for x in got if x[2] == "G3.t1"]
        b = [x for x in got if x[2] == "G4.t1"]
        assert len(a) == 1 and len(b) == 1, got
        assert a[0][4] == "300M", a                                                  # A's own clip rescued into E1
        assert b[0][4] == "225M" and b[0][3] == 200 - 25, b                          # B's clip rescued with A's bases into F1
    fixture("first_record_seq", ann, recs, ["--lr"], chk, genome=[bytes(g)], doc="""
first_record_seq: (--lr -S) two records of one read name; the second is reverse-flagged and has no sequence ('*').  Its 25-base
left clip is rescued by ksw2 against the upstream exon using the FIRST record's sequence as is (src/core.cpp:353-378).""")

    # 4 -- an I directly between two Ns suppresses closing the (empty) exon (gclib/GSam.cpp:245-261)
    ann = synth.Annotation(["chr1"], [300000], [tx("G5.t1", "+", [(40000, 40049), (40950, 41100)])])
    recs = [rec("i1", 0, 40000, [("M", 50), ("N", 500), ("I", 3), ("N", 400), ("M", 47)], tags=[("NH", "C", 1), ("XS", "A", "+")])]

    def chk(got):
        assert len(got) == 1 and got[0][4] == "50M3I47M" and got[0][3] == 0, got
    fixture("ins_between_introns", ann, recs, [], chk, doc="""
ins_between_introns: CIGAR 50M500N3I400N47M -- the I between the two Ns keeps setupCoordinates from closing an empty exon
(gclib/GSam.cpp:245-261): two read exons, projected CIGAR 50M3I47M.""")

    # 5 -- uint8_t exon_id wraps at 256 exons (src/bramble.cpp:170, src/evaluate.cpp:131): skipping 256 guide exons looks like a
    #      junction between consecutive exons (gap == 1), skipping 255 does not
    ex = [(50000 + 200 * k, 50000 + 200 * k + 99) for k in range(300)]
    ann = synth.Annotation(["chr1"], [300000], [tx("G6.t1", "+", ex)])
    xs = [("NH", "C", 1), ("XS", "A", "+")]
    recs = [rec("w_adjacent", 0, ex[255][1] - 49, [("M", 50), ("N", 100), ("M", 50)], tags=xs),                       # exon 255 -> 256: id 255 -> 0
            rec("w_skip256", 0, ex[10][1] - 49, [("M", 50), ("N", ex[267][0] - ex[10][1] - 1), ("M", 50)], tags=xs),      # exon 10 -> 267: (267-10) mod 256 == 1
            rec("w_skip255", 0, ex[10][1] - 49, [("M", 50), ("N", ex[266][0] - ex[10][1] - 1), ("M", 50)], tags=xs)]      # exon 10 -> 266: gap 0 -> eliminated

    def chk(got):
        assert sorted(x[0] for x in got) == ["w_adjacent", "w_skip256"], got
        assert all(x[4] == "100M" for x in got)
    fixture("exon_id_wrap_256", ann, recs, [], chk, doc="""
exon_id_wrap_256: a 300-exon transcript; exon_id is a uint8_t (src/bramble.cpp:170).  A junction exon 255 -> 256 is accepted
(255 -> 0, difference 1 mod 256); a read that jumps from exon 10 to exon 267 is ALSO accepted ((267 - 10) mod 256 == 1,
src/evaluate.cpp:131-140); exon 10 -> 266 (difference 0) is eliminated.""")

    # 6 -- src/g2t.cpp:204 `status == FIRST_EXON || MIDDLE_EXON` is always true: on a '-' strand query a right overhang is always
    #      tested against max_junc_ins (0 for short reads), never against max_clip (5)
    ann = synth.Annotation(["chr1"], [300000], [tx("G7.t1", "+", [(150000, 150300)]), tx("G8.t1", "-", [(150000, 150300)])])
    un = [("NH", "C", 1)]
    recs = [rec("o_right3", 0, 150244, [("M", 60)], tags=un),       # overhangs the exon end by 3 <= max_clip: '+' accepts, '-' rejects
            rec("o_left3", 0, 149997, [("M", 60)], tags=un),        # overhangs the exon start by 3: both strands accept
            rec("o_inside", 0, 150100, [("M", 60)], tags=un)]

    def chk(got):
        by = {}
        for x in got:
            by.setdefault(x[0], []).append(x[2])
        assert sorted(by["o_right3"]) == ["G7.t1"], by
        assert sorted(by["o_left3"]) == ["G7.t1", "G8.t1"], by
        assert sorted(by["o_inside"]) == ["G7.t1", "G8.t1"], by
    fixture("minus_right_overhang", ann, recs, [], chk, doc="""
minus_right_overhang: identical single-exon transcripts on '+' and '-'; unstranded reads search both.  A 3-base overhang past the
exon END passes max_clip on '+' but is tested against max_junc_ins == 0 on '-' (src/g2t.cpp:204, the always-true `|| MIDDLE_EXON`);
a 3-base overhang before the exon START passes on both.""")

    # 7 -- first_match_idx / last_match_idx are COUNTERS of guide-bearing segments compared with the segment index
    #      (src/evaluate.cpp:1010-1011,1053-1054,1075-1076): behind an INS_EXON segment the MIDDLE exon at k == #guide-1 is
    #      treated as the last match and its right_gap emits no D
    A_, B_, C_ = (200000, 200299), (201000, 201199), (202000, 202299)
    ann = synth.Annotation(["chr1"], [300000], [tx("G9.t1", "+", [A_, B_, C_])])
    lt = [("NM", "C", 0), ("AS", "i", 400), ("ts", "A", "+"), ("NH", "C", 1)]

    def spliced(pos1, blocks):
        cg = []
        p = pos1
        for i, (s, e) in enumerate(blocks):
            if i:
                cg.append(("N", s - p))
            cg.append(("M", e - s + 1))
            p = e + 1
        return cg
    with_ins = [(200100, 200299), (200600, 200619), (201000, 201189), (202000, 202199)]      # spurious 20-base exon in the intron; B ends 10 short
    without = [(200100, 200299), (201000, 201189), (202000, 202199)]
    recs = [rec("c_ins", 0, 200100, spliced(200100, with_ins), tags=lt), rec("c_plain", 0, 200100, spliced(200100, without), tags=lt)]

    def chk(got):
        by = {x[0]: x[4] for x in got}
        assert "D" not in by["c_ins"] and "20I" in by["c_ins"], by          # the 10-base right_gap of exon B emits no D behind the INS segment
        assert "10D" in by["c_plain"], by                                   # ... and does without it
    fixture("match_idx_counters", ann, recs, ["--lr"], chk, doc="""
match_idx_counters: (--lr) read exon B ends 10 bases before its guide exon ends (right_gap 10 <= max_junc_gap).  Without anything
else the ideal CIGAR gets a 10D there.  With a spurious 20-base exon (an INS_EXON segment) earlier in the read, last_match_idx
(a count of guide-bearing segments, src/evaluate.cpp:1010-1011) equals B's segment index, B is taken for the last match and the D
is not emitted (:1075-1076).""")

    # 8 -- aux carry-through: NH re-set on the shared source record, XS dropped, HI re-appended; every other tag kept in place and order
    ann = synth.Annotation(["chr1"], [300000], [tx("G10.t1", "+", [(250000, 250400)]), tx("G10.t2", "+", [(250000, 250400), (251000, 251200)]),
                                               tx("G10.t3", "+", [(249900, 250400)])])
    odd = [("XS", "A", "+"), ("NM", "i", 1), ("HI", "i", 5), ("ZZ", "Z", "hello"), ("NH", "C", 7), ("AS", "s", 99), ("BQ", "f", 0.5)]
    recs = [rec("t1", 0, 250100, [("M", 100)], tags=odd)]

    def chk(got):
        assert len(got) == 3, got
        for x in got:
            aux = x[5]
            assert aux.startswith(b"NMi\x01\x00\x00\x00ZZZhello\x00ASs\x63\x00BQf") and b"XSA" not in aux, aux
            assert aux[-14:-7] == b"NHi\x03\x00\x00\x00" and aux[-7:-4] == b"HIi", aux
    fixture("aux_carry_through", ann, recs, [], chk, doc="""
aux_carry_through: one read on three isoforms; input tags in odd order and types.  Output aux = input minus NH, XS and HI, in input
order, then NH:i (record count of the name) and HI:i (src/core.cpp:115-124, src/bam.cpp:590-620).""")

    # 9 -- record bytes: reverse complement of an odd-length read, missing qualities, array / hex / string tags kept, duplicate and
    #      oddly typed NH / HI / XS, mate fields of pairs on a '-' transcript and across two transcripts
    ann = synth.Annotation(["chr1"], [300000], [tx("G11.t1", "-", [(260000, 260500), (261000, 261400)]), tx("G11.t2", "+", [(270000, 270600)]),
                                               tx("G11.t3", "+", [(270000, 270600), (271000, 271200)])])
    zoo = [("NH", "s", 300), ("XA", "B", ("s", [-3, 7, 30000])), ("XS", "A", "-"), ("HI", "S", 65535), ("MD", "Z", "101"), ("NH", "i", 9), ("XH", "H", "1AE301"),
           ("XS", "A", "+"), ("XB", "B", ("C", [])), ("XF", "f", 1.25), ("HI", "c", -1)]
    r_odd = rec("z_odd", 16, 260100, [("M", 101)], tags=zoo)
    r_noq = rec("z_noq", 0, 260150, [("M", 75)], tags=[("XS", "A", "-")]); r_noq["qual"] = None
    r_spl = rec("z_spl", 16, 260451, [("M", 50), ("N", 499), ("M", 51)], tags=zoo[2:])
    recs = [r_odd, r_noq, r_spl,
            rec("z_pair", 99, 260050, [("M", 99)], 260300, tags=zoo[:4]), rec("z_pair", 147, 260300, [("M", 100)], 260050, tags=[("XS", "A", "-")] + zoo[4:7]),
            rec("z_two", 99, 270100, [("M", 100)], 270350, tags=[]), rec("z_two", 147, 270350, [("M", 100)], 270100, tags=[("NH", "C", 1)])]

    def chk(got):
        by = {}
        for x in got:
            by.setdefault(x[0], []).append(x)
        assert len(by["z_odd"]) == 1 and len(by["z_noq"]) == 1 and len(by["z_spl"]) == 1 and len(by["z_pair"]) == 2 and len(by["z_two"]) == 4, {k: len(v) for k, v in by.items()}
        aux = by["z_odd"][0][5]
        assert aux.startswith(b"XABs\x03\x00\x00\x00") and b"MDZ101\x00NHi\x09\x00\x00\x00XHH1AE301\x00XSA+XBBC\x00\x00\x00\x00XFf" in aux, aux    # first NH / XS / HI only
        assert aux.count(b"HIc\xff") == 1 and aux.endswith(b"NHi\x01\x00\x00\x00HIi\x01\x00\x00\x00"), aux
    fixture("record_bytes_zoo", ann, recs, [], chk, doc="""
record_bytes_zoo: what write_to_bam does to the bytes of a record (src/core.cpp:96-212, src/bam.cpp:475-702): an odd-length read and a
spliced read on a '-' transcript (sequence reverse-complemented nibble by nibble, qualities reversed, CIGAR reversed, flag 16 toggled), a read
without qualities (0xff kept as is), B / H / Z / f tags carried through, only the FIRST NH, XS and HI dropped whatever their type, pairs
on a '-' transcript (MREVERSE from the own strand, signed TLEN) and a pair that lands on two transcripts twice (4 records).""")

    # 10 -- the same for long reads: ts dropped instead of XS, AS rewritten from its value whatever the integer type
    ann = synth.Annotation(["chr1"], [300000], [tx("G12.t1", "-", [(280000, 280900)]), tx("G12.t2", "+", [(281500, 282400)])])
    recs = []
    for k, (ty, val) in enumerate([("c", -5), ("C", 200), ("s", -300), ("S", 60000), ("i", 123456), ("I", 77)]):
        recs.append(rec("l%d" % k, 0 if k % 2 else 16, (280100 if k % 3 else 281600) + 10 * k, [("M", 301 + k)],
                        tags=[("ts", "A", "+"), ("XS", "A", "-"), ("AS", ty, val), ("NM", "i", k), ("ts", "A", "-"), ("AS", "i", 1)]))
    recs.append(rec("l_noas", 0, 280300, [("M", 250)], tags=[("NM", "i", 3)]))

    def chk(got):
        assert len(got) == 7, got
        for x in got:
            aux = x[5]
            if x[0] != "l_noas":
                assert aux.startswith(b"XSA-NMi") and aux.count(b"tsA") == 1 and aux.count(b"ASi") == 2, aux      # first ts and first AS gone, the second ones stay
            assert aux[-21:-18] == b"NHi" and aux[-14:-11] == b"ASi" and aux[-7:-4] == b"HIi", aux
    fixture("record_bytes_zoo_lr", ann, recs, ["--lr"], chk, doc="""
record_bytes_zoo_lr: (--lr) the first ts is dropped (XS stays), the first AS is read as an integer of whatever type it has, dropped, and a new
AS:i = (AS + clip score) * similarity is appended between NH:i and HI:i (set_as_tag, src/bam.cpp:622-634); a read without AS gets AS:i from 0.""")


if __name__ == "__main__":
    main()
