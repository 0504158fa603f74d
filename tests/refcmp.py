"""TEST INFRASTRUCTURE: turn engine/oracle output rows into the BAM records the reference would
write (record assembly of src/core.cpp:96-212 + flush() :237-332, src/bam.cpp:531-634), and parse the
reference binary's output BAM into the same tuples, so the two can be compared name by name.
The SECONDARY flag is excluded when a name has a tie for the best score (the winner is picked by a
libstdc++ RNG on the host; that part is covered by the byte-exact CLI test)."""
import struct
from collections import defaultdict

import numpy as np

from bramble_b200 import capi

OPS = "MIDNSHP=XB??????"


def cigar_str(words):
    return "".join(f"{int(w) >> 4}{OPS[int(w) & 15]}" for w in words) or "*"


def get_mapq(nh, long_reads):
    if not long_reads:
        return 255 if nh == 1 else 3 if nh == 2 else 1 if nh in (3, 4) else 0
    return 0 if nh > 1 else 3


def _old_as(rec):
    for t, ty, v in rec.get("tags", []):
        if t == "AS":
            return int(v)
    return 0


def expected_records(batches_rows, recs, long_reads):
    """-> dict name -> list of tuples in HI order.  batches_rows: [(Batch, rows dict)]."""
    by_name = defaultdict(list)
    for batch, rows in batches_rows:
        n = rows["n_rec"]
        goff = rows["group_rec_off"]
        for g in range(len(goff) - 1):
            lo, hi = int(goff[g]), int(goff[g + 1])
            if hi == lo:
                continue
            nh = hi - lo
            mapq = get_mapq(nh, long_reads)
            # best pair score bookkeeping (flush())
            units = []
            r = lo
            while r < hi:
                paired = bool(rows["flags"][r] & capi.REC_PAIRED)
                sc = rows["sim_score"][r]
                if paired:
                    sc = max(sc, rows["sim_score"][r + 1])
                units.append((r, paired, sc))
                r += 2 if paired else 1
            best = max(u[2] for u in units)
            nbest = sum(1 for u in units if u[2] == best)
            for (r0, paired, sc) in units:
                for k in range(2 if paired else 1):
                    r = r0 + k
                    a = int(rows["src_aln"][r])
                    rec = recs[int(batch.src[a])]
                    flag = rec["flag"]
                    strand = chr(rows["tx_strand"][r])
                    if strand == "-":
                        flag ^= 16
                    tid, pos = int(rows["tid"][r]), int(rows["pos"][r])
                    lq = len(rec["seq"] or b"")
                    if not paired:
                        flag &= ~(1 | 2 | 32)
                        mtid, mpos, isize = -1, -1, 0
                    else:
                        o = r0 + (1 - k)
                        flag |= 1
                        if strand == "-":
                            flag |= 32
                        opos, otid = int(rows["pos"][o]), int(rows["tid"][o])
                        if rows["flags"][r] & capi.REC_SAMETX:
                            mtid, mpos = tid, opos
                            flag |= 2
                            isize = (opos + lq) - pos if pos <= opos else -((pos + lq) - opos)
                        else:
                            mtid, mpos, isize = otid, opos, 0
                            flag &= ~2
                    prim = None
                    if nbest == 1:
                        prim = (sc == best)
                        flag = (flag & ~256) if prim else (flag | 256)
                    else:
                        flag &= ~256
                    asv = None
                    if long_reads:
                        asv = int((float(_old_as(rec) + int(rows["clip_score"][r]))) * float(rows["sim_score"][r]))
                    name = rec["name"]
                    by_name[name].append((tid, pos, flag, mapq, cigar_str(rows["cigars"][r]), mtid, mpos, isize, nh,
                                          r - lo + 1, asv, nbest == 1))
    return by_name


def _aux_ints(aux):
    out = {}
    p = 0
    n = len(aux)
    sz = {"c": 1, "C": 1, "s": 2, "S": 2, "i": 4, "I": 4, "f": 4, "A": 1}
    fmt = {"c": "b", "C": "B", "s": "h", "S": "H", "i": "i", "I": "I"}
    while p + 3 <= n:
        tag = aux[p:p + 2].decode()
        ty = chr(aux[p + 2])
        p += 3
        if ty in fmt:
            out[tag] = struct.unpack_from("<" + fmt[ty], aux, p)[0]
            p += sz[ty]
        elif ty in ("A", "f"):
            p += sz[ty]
        elif ty in ("Z", "H"):
            while aux[p] != 0:
                p += 1
            p += 1
        elif ty == "B":
            sub = chr(aux[p])
            cnt = struct.unpack_from("<I", aux, p + 1)[0]
            p += 5 + cnt * sz[sub]
        else:
            raise ValueError(ty)
    return out


def parse_output_bam(raw_recs, long_reads):
    """raw BAM record bodies (synth.read_bam) -> dict name -> list of tuples sorted by HI."""
    by_name = defaultdict(list)
    for rec in raw_recs:
        tid, pos, bmq, fnc, lq, mtid, mpos, isz = struct.unpack_from("<iiIIiiii", rec, 0)
        lname = bmq & 0xff
        ncig = fnc & 0xffff
        flag = fnc >> 16
        name = rec[32:32 + lname - 1].decode()
        cig = struct.unpack_from(f"<{ncig}I", rec, 32 + lname)
        aux = rec[32 + lname + 4 * ncig + (lq + 1) // 2 + lq:]
        tags = _aux_ints(aux)
        by_name[name].append([tid, pos, flag, (bmq >> 8) & 0xff, cigar_str(cig), mtid, mpos, isz, tags.get("NH"), tags.get("HI"),
                              tags.get("AS") if long_reads else None])
    for name, lst in by_name.items():
        lst.sort(key=lambda t: t[9])
    return by_name


def compare(expected, got):
    """expected from expected_records, got from parse_output_bam."""
    assert set(expected) == set(got), f"names differ: only expected {sorted(set(expected) - set(got))[:5]}, only got {sorted(set(got) - set(expected))[:5]}"
    for name, exp in expected.items():
        g = got[name]
        assert len(exp) == len(g), f"{name}: {len(exp)} vs {len(g)} records\n{exp}\n{g}"
        for e, r in zip(exp, g):
            unique_best = e[11]
            ef = e[2]
            rf = r[2] if unique_best else (r[2] & ~256)
            et = (e[0], e[1], ef) + tuple(e[3:11])
            rt = (r[0], r[1], rf) + tuple(r[3:11])
            assert et == rt, f"{name}:\n exp {et}\n got {rt}"
        if not exp[0][11]:
            # tie: exactly one unit is primary
            prim = [r for r in g if not (r[2] & 256)]
            assert 1 <= len(prim) <= 2
