"""Host-side packing for the projection engine: annotation -> flattened g2t interval table,
alignments -> structure-of-arrays bundles (`brb_index_desc` / `brb_batch_in` of include/bramble_b200.h).

This is the Python mirror of the ingest half of the reference (what happens BEFORE convert_reads):
  * index:   build_g2t_tree            reference src/bramble.cpp:132-211, src/g2t.cpp:34-62
  * exons:   GSamRecord::setupCoordinates + process_exons   gclib/GSam.cpp:197-291, src/bramble.cpp:246-255
  * strand:  get_strand / spliceStrand  src/bramble.cpp:213-244, gclib/GSam.cpp:338-349
  * mates:   process_pairs              src/bramble.cpp:272-311
  * bundles: process_reads              src/bramble.cpp:329-435 (cut at >=100000 alignments on a name change)
It does no projection arithmetic; that lives in the CUDA engine (csrc/).
"""
import ctypes as C

import numpy as np

from . import capi

OPS = "MIDNSHP=X"
_OPI = {c: i for i, c in enumerate(OPS)}
_CODE = np.full(256, 4, dtype=np.uint8)
for _i, _c in enumerate(b"ACGT"):
    _CODE[_c] = _i


class Index:
    """Flattened interval table + optional genome; owns the numpy arrays the descriptor points to."""

    def __init__(self):
        self.desc = capi.IndexDesc()
        self.arrays = {}
        self.tid_names = []
        self.tid_len = []
        self.tid_strand = []
        self.refid_of = {}

    @property
    def n_records(self):
        return int(self.arrays["st"].shape[0])


def sort_perm_by_start(st):
    """Permutation cgranges' IITree::index() applies: libstdc++ std::sort (unstable) on `st` only.
    Done by the host half of the engine library so that tie order is the real introsort's."""
    st = np.ascontiguousarray(st, dtype=np.int32)
    perm = np.empty(st.shape[0], dtype=np.uint32)
    if st.shape[0]:
        capi.check(capi.lib().brb_sort_by_start(st.shape[0], capi.ptr(st), capi.ptr(perm)), "brb_sort_by_start")
    return perm


def build_index(ann, gtf_order=None, genome=None, sorter=sort_perm_by_start):
    """ann: synth.Annotation-like (chroms, tx[dict(id, chrom, strand, exons 1-based inclusive ascending)]).
    gtf_order: order in which transcripts appear in the GTF file (defines refid numbering, the
    gclib gseq_id = first appearance); tids follow the sorted order (chrom name, start, end, id) of
    gclib gfo_cmpByLoc, gff.cpp:75-91."""
    ix = Index()
    order = list(range(len(ann.tx))) if gtf_order is None else list(gtf_order)
    for i in order:
        c = ann.chroms[ann.tx[i]["chrom"]]
        if c not in ix.refid_of:
            ix.refid_of[c] = len(ix.refid_of)
    n_ref = len(ix.refid_of)
    srt = sorted(range(len(ann.tx)), key=lambda i: (ann.chroms[ann.tx[i]["chrom"]].encode(), ann.tx[i]["exons"][0][0],
                                                    ann.tx[i]["exons"][-1][1], ann.tx[i]["id"].encode()))
    tables = [[] for _ in range(2 * n_ref)]
    for tid, i in enumerate(srt):
        t = ann.tx[i]
        ix.tid_names.append(t["id"])
        ix.tid_strand.append(t["strand"])
        ex = t["exons"]
        n = len(ex)
        iv = []
        pos_start = 0
        for k in range(n):
            idx = n - k - 1 if t["strand"] == "-" else k
            s, e = ex[idx][0], ex[idx][1] + 1
            iv.append([s, e, idx & 0xff, pos_start])
            pos_start += e - s
        ix.tid_len.append(pos_start)
        if t["strand"] not in "+-":
            continue
        tb = tables[2 * ix.refid_of[ann.chroms[t["chrom"]]] + (t["strand"] == "-")]
        for k in range(n):
            s, e, idx, ps = iv[k]
            pv = (iv[k - 1][0], iv[k - 1][1]) if k > 0 else (0, 0)
            nx = (iv[k + 1][0], iv[k + 1][1]) if k < n - 1 else (0, 0)
            tb.append((s, e, tid, idx, ps, pv[0], pv[1], nx[0], nx[1], (1 if k > 0 else 0) | (2 if k < n - 1 else 0), pos_start))
    cols = [[] for _ in range(11)]
    table_off = [0]
    for tb in tables:
        if tb:
            a = np.array(tb, dtype=np.int64)
            perm = sorter(a[:, 0].astype(np.int32))
            a = a[perm]
            for c in range(11):
                cols[c].append(a[:, c])
        table_off.append(table_off[-1] + len(tb))
    names = ["st", "en", "tid", "exon_id", "pos_start", "prev_st", "prev_en", "next_st", "next_en", "has_prev_next", "transcript_len"]
    dts = [np.int32, np.int32, np.uint32, np.uint8, np.uint32, np.uint32, np.uint32, np.uint32, np.uint32, np.uint8, np.uint32]
    A = ix.arrays
    for nme, dt, col in zip(names, dts, cols):
        A[nme] = np.ascontiguousarray(np.concatenate(col).astype(dt)) if col else np.zeros(0, dtype=dt)
    A["table_off"] = np.array(table_off, dtype=np.uint64)
    d = ix.desc
    d.n_tables = 2 * n_ref
    d.table_off = capi.ptr(A["table_off"])
    for nme in names:
        setattr(d, nme, capi.ptr(A[nme]))
    d.n_chrom = 0
    d.chrom_off = None
    d.genome_codes = None
    if genome is not None:
        # per-refid base codes; genome[] is indexed like ann.chroms
        by_ref = [None] * n_ref
        for ci, cname in enumerate(ann.chroms):
            if cname in ix.refid_of:
                by_ref[ix.refid_of[cname]] = genome[ci]
        offs = [0]
        parts = []
        for g in by_ref:
            codes = _CODE[np.frombuffer(g.upper(), dtype=np.uint8)] if g is not None else np.zeros(0, np.uint8)
            parts.append(codes)
            offs.append(offs[-1] + len(codes))
        A["genome_codes"] = np.ascontiguousarray(np.concatenate(parts)) if parts else np.zeros(0, np.uint8)
        A["chrom_off"] = np.array(offs, dtype=np.uint64)
        d.n_chrom = n_ref
        d.chrom_off = capi.ptr(A["chrom_off"])
        d.genome_codes = capi.ptr(A["genome_codes"])
    return ix


# ----------------------------------------------------------------------------- alignments
def cigar_words(cig):
    return [(l << 4) | _OPI[o] for o, l in cig]


def read_exons(pos0, words):
    """GSamRecord::setupCoordinates (gclib/GSam.cpp:197-291) followed by process_exons' end++
    (src/bramble.cpp:246-255): list of [start, end) 1-based half-open read exons."""
    l = 0
    exstart = pos0
    ex_s = ex_e = 0
    started = intron = ins = False
    out = []
    for w in words:
        op, ln = w & 0xf, w >> 4
        if op in (0, 7, 8):
            started = True
            l += ln
            intron = False
            ins = False
        elif op == 2:
            l += ln
            ins = False
        elif op == 1:
            ins = True
        elif op == 3:
            if not started:
                continue
            if (not ins) or (not intron):
                ex_e = pos0 + l
                ex_s = exstart + 1
                out.append((ex_s, ex_e + 1))
            l += ln
            exstart = pos0 + l
            intron = True
        elif op in (4, 5):
            ins = False
    if not intron:
        ex_s = exstart + 1
        ex_e = pos0 + l
        out.append((ex_s, ex_e + 1))
    if ex_e == 0:
        raise ValueError("invalid CIGAR record")   # GError in the reference
    return out


def _tag(rec, name):
    for t, ty, v in rec.get("tags", []):
        if t == name:
            return ty, v
    return None


def read_strand(rec, fr=False, rf=False):
    """spliceStrand (XS, else minimap2 ts flipped by FREVERSE) then the --fr/--rf rule."""
    c = 0
    t = _tag(rec, "XS")
    if t and t[0] in "AZ" and t[1]:
        c = t[1][0]
    if c == 0:
        t = _tag(rec, "ts")
        m = t[1][0] if (t and t[0] in "AZ" and t[1]) else 0
        if m in ("+", "-"):
            c = ({"+": "-", "-": "+"}[m]) if rec["flag"] & 16 else m
    strand = c if c in ("+", "-") else "."
    if strand == "." and (fr or rf):
        rev = bool(rec["flag"] & 16)
        hit = (rf and rev) or (fr and not rev)
        if rec["flag"] & 1:
            order = 1 if rec["flag"] & 0x40 else (2 if rec["flag"] & 0x80 else 0)
            if order == 1:
                strand = "-" if hit else "+"
            else:
                strand = "+" if hit else "-"
        else:
            strand = "-" if hit else "+"
    return strand


class Batch:
    """One bundle as SoA; owns the arrays `desc` points to.  `src` maps alignment -> input record index."""

    def __init__(self):
        self.desc = capi.BatchIn()
        self.arrays = {}
        self.src = None
        self.src_offset = 0
        self.names = []

    def finish(self, refid, strand, flag, exon_off, exon_st, exon_en, cigar_off, cigar, mate, group_off, seq_off, seq_codes, src, names):
        A = self.arrays
        A["refid"] = np.array(refid, dtype=np.int32)
        A["strand"] = np.array(strand, dtype=np.int8)
        A["exon_off"] = np.array(exon_off, dtype=np.uint32)
        A["exon_st"] = np.array(exon_st, dtype=np.uint32)
        A["exon_en"] = np.array(exon_en, dtype=np.uint32)
        A["cigar_off"] = np.array(cigar_off, dtype=np.uint32)
        A["cigar"] = np.array(cigar, dtype=np.uint32)
        A["mate_idx"] = np.array(mate, dtype=np.int32)
        A["group_off"] = np.array(group_off, dtype=np.uint32)
        self.src = np.array(src, dtype=np.int64)
        self.names = names
        d = self.desc
        d.n_aln = len(refid)
        d.n_groups = len(group_off) - 1
        for k in ("group_off", "refid", "strand", "exon_off", "exon_st", "exon_en", "cigar_off", "cigar", "mate_idx"):
            setattr(d, k, capi.ptr(A[k]))
        if seq_off is not None:
            A["group_seq_off"] = np.array(seq_off, dtype=np.uint64)
            A["seq_codes"] = np.ascontiguousarray(np.concatenate(seq_codes)) if seq_codes else np.zeros(0, np.uint8)
            if A["seq_codes"].shape[0] == 0:
                A["seq_codes"] = np.zeros(1, np.uint8)
            d.group_seq_off = capi.ptr(A["group_seq_off"])
            d.seq_codes = capi.ptr(A["seq_codes"])
        else:
            d.group_seq_off = None
            d.seq_codes = None
        return self

    def attach_names(self, group_names):
        """Read names per group (brb_batch_in::group_name_off / name_bytes): what BRB_OPT_FLUSH_TAGS hashes for the primary tie-break."""
        enc = [n.encode() if isinstance(n, str) else bytes(n) for n in group_names]
        assert len(enc) == int(self.desc.n_groups)
        A = self.arrays
        A["group_name_off"] = np.concatenate([[0], np.cumsum([len(x) for x in enc])]).astype(np.uint64)
        A["name_bytes"] = np.frombuffer(b"".join(enc) + b"\0", dtype=np.uint8).copy()
        self.desc.group_name_off = capi.ptr(A["group_name_off"]); self.desc.name_bytes = capi.ptr(A["name_bytes"])
        return self

    def compact(self, exons=True):
        """The same bundle with the compact input columns of brb_batch_in (fewer bytes host -> device): 16-bit CIGAR word counts, 16-bit
        mate deltas and refids, and -- `exons` -- the 0-based position instead of the read-exon arrays (the engine derives the exons
        from position + CIGAR as the reference's ingest does).  The columns they replace are left out (NULL)."""
        A = self.arrays
        n = int(self.desc.n_aln)
        nb = Batch()
        B = nb.arrays
        for k in ("group_off", "strand", "cigar", "group_seq_off", "seq_codes", "group_name_off", "name_bytes"):
            if k in A:
                B[k] = A[k]
        co = A["cigar_off"].astype(np.int64)[:n + 1]
        cnt = np.diff(co)
        mate = A["mate_idx"].astype(np.int64)[:n]
        delta = np.where(mate >= 0, mate - np.arange(n), 0)
        refid = A["refid"].astype(np.int64)[:n]
        if n and (cnt.max() > 65535 or np.abs(delta).max() > 32767 or refid.max() > 32767 or refid.min() < -32768):
            raise ValueError("bundle does not fit the compact columns")
        B["cigar_cnt16"] = np.ascontiguousarray(cnt, dtype=np.uint16) if n else np.zeros(1, np.uint16)
        B["mate_delta16"] = np.ascontiguousarray(delta, dtype=np.int16) if n else np.zeros(1, np.int16)
        B["refid16"] = np.ascontiguousarray(refid, dtype=np.int16) if n else np.zeros(1, np.int16)
        if exons:
            eo = A["exon_off"].astype(np.int64)[:n]
            B["pos0"] = (A["exon_st"].astype(np.int64)[eo] - 1).astype(np.int32) if n else np.zeros(1, np.int32)      # the first read exon starts at pos0 + 1
        else:
            for k in ("exon_off", "exon_st", "exon_en"):
                B[k] = A[k]
        nb.src, nb.src_offset, nb.names = self.src, self.src_offset, self.names
        d = nb.desc
        d.n_aln, d.n_groups = self.desc.n_aln, self.desc.n_groups
        d.n_cigar_total = int(co[n]) if n else 0
        for k, a in B.items():
            if hasattr(d, k):
                setattr(d, k, capi.ptr(a))
        # (the decoded columns stay reachable for code that reads them on the host: row unpacking needs mate_idx)
        nb.arrays_full = A
        return nb

    def slice_groups(self, g0, g1):
        """Read-name groups [g0, g1) as a bundle of their own (mates never leave their group, src/bramble.cpp:272-311)."""
        A = self.arrays
        go = A["group_off"].astype(np.int64)
        a0, a1 = int(go[g0]), int(go[g1])
        eo, co = A["exon_off"].astype(np.int64), A["cigar_off"].astype(np.int64)
        mate = A["mate_idx"][a0:a1].astype(np.int64)
        nb = Batch()
        B = nb.arrays
        B["refid"] = np.ascontiguousarray(A["refid"][a0:a1]); B["strand"] = np.ascontiguousarray(A["strand"][a0:a1])
        B["exon_off"] = (eo[a0:a1 + 1] - eo[a0]).astype(np.uint32)
        B["exon_st"] = np.ascontiguousarray(A["exon_st"][eo[a0]:eo[a1]]); B["exon_en"] = np.ascontiguousarray(A["exon_en"][eo[a0]:eo[a1]])
        B["cigar_off"] = (co[a0:a1 + 1] - co[a0]).astype(np.uint32)
        B["cigar"] = np.ascontiguousarray(A["cigar"][co[a0]:co[a1]])
        B["mate_idx"] = np.where(mate >= 0, mate - a0, -1).astype(np.int32)
        B["group_off"] = (go[g0:g1 + 1] - a0).astype(np.uint32)
        nb.src_offset = a0                                   # alignment index of the slice's first alignment in the parent
        nb.src = self.src[a0:a1] if self.src is not None and len(self.src) else None
        nb.names = self.names[g0:g1] if self.names else []
        d = nb.desc
        d.n_aln, d.n_groups = a1 - a0, g1 - g0
        for k in ("group_off", "refid", "strand", "exon_off", "exon_st", "exon_en", "cigar_off", "cigar", "mate_idx"):
            if B[k].size == 0:
                B[k] = np.zeros(1, B[k].dtype)
            setattr(d, k, capi.ptr(B[k]))
        if "group_seq_off" in A:
            so = A["group_seq_off"].astype(np.int64)
            B["group_seq_off"] = (so[g0:g1 + 1] - so[g0]).astype(np.uint64)
            B["seq_codes"] = np.ascontiguousarray(A["seq_codes"][so[g0]:so[g1]])
            if B["seq_codes"].size == 0:
                B["seq_codes"] = np.zeros(1, np.uint8)
            d.group_seq_off = capi.ptr(B["group_seq_off"]); d.seq_codes = capi.ptr(B["seq_codes"])
        else:
            d.group_seq_off = None; d.seq_codes = None
        return nb


def concat_batches(parts):
    """Bundles (their `arrays` dicts) laid end to end as one bundle: CSR offsets and mate links shifted."""
    nb = Batch()
    B = nb.arrays
    n_al = np.array([0] + [p["refid"].shape[0] for p in parts], dtype=np.int64).cumsum()
    for k in ("refid", "strand", "exon_st", "exon_en", "cigar"):
        B[k] = np.ascontiguousarray(np.concatenate([p[k] for p in parts]))
    for off, dat in (("exon_off", "exon_st"), ("cigar_off", "cigar"), ("group_off", "refid")):
        base = np.array([0] + [p[dat].shape[0] for p in parts], dtype=np.int64).cumsum()
        B[off] = np.concatenate([p[off][(1 if i else 0):].astype(np.int64) + base[i] for i, p in enumerate(parts)]).astype(np.uint32)
    B["mate_idx"] = np.concatenate([np.where(p["mate_idx"] >= 0, p["mate_idx"].astype(np.int64) + n_al[i], -1) for i, p in enumerate(parts)]).astype(np.int32)
    d = nb.desc
    d.n_aln, d.n_groups = int(n_al[-1]), int(B["group_off"].shape[0] - 1)
    for k in ("group_off", "refid", "strand", "exon_off", "exon_st", "exon_en", "cigar_off", "cigar", "mate_idx"):
        setattr(d, k, capi.ptr(B[k]))
    if all("group_seq_off" in p for p in parts):
        base = np.array([0] + [int(p["group_seq_off"][-1]) for p in parts], dtype=np.int64).cumsum()
        B["group_seq_off"] = np.concatenate([p["group_seq_off"][(1 if i else 0):].astype(np.int64) + base[i] for i, p in enumerate(parts)]).astype(np.uint64)
        B["seq_codes"] = np.ascontiguousarray(np.concatenate([p["seq_codes"][:int(p["group_seq_off"][-1])] for p in parts] + [np.zeros(1, np.uint8)]))
        d.group_seq_off = capi.ptr(B["group_seq_off"]); d.seq_codes = capi.ptr(B["seq_codes"])
    else:
        d.group_seq_off = None; d.seq_codes = None
    nb.src = np.arange(d.n_aln, dtype=np.int64)
    return nb


_NT16 = "=ACMGRSVTWYHKDBN"


def _seq_codes(seq):
    """Group-shared sequence as the rescue sees it: BAM 4-bit -> seq_nt16_str char -> ksw2 code
    (src/evaluate.cpp:299-305, :320-330).  A byte string of IUPAC letters goes through the same table."""
    b = np.frombuffer(bytes(seq), dtype=np.uint8)
    # BAM stores letters outside =ACMGRSVTWYHKDBN as N; all of those are code 4 anyway
    return _CODE[b].copy()


def pack_batches(recs, index, ann_chroms, long_reads=False, fr=False, rf=False, use_fasta=False, chunk_size=100000):
    """Split BAM-ordered record dicts into bundles exactly like process_reads and pack each one."""
    out = []
    cur = None

    def new_state():
        return dict(refid=[], strand=[], flag=[], exon_off=[0], exon_st=[], exon_en=[], cigar_off=[0], cigar=[], mate=[],
                    group_off=[0], src=[], names=[], hashread={}, seqs=[], start=[], group_seq=None)

    def close(st):
        if not st["refid"]:
            return
        n = len(st["refid"])
        st["group_off"].append(n)
        seq_off = seq_codes = None
        if use_fasta:
            seq_off = [0]
            seq_codes = []
            for g in range(len(st["group_off"]) - 1):
                s = np.zeros(0, np.uint8)
                for a in range(st["group_off"][g], st["group_off"][g + 1]):
                    if len(st["seqs"][a]) > 0:
                        s = _seq_codes(st["seqs"][a])
                        break
                seq_codes.append(s)
                seq_off.append(seq_off[-1] + len(s))
        b = Batch().finish(st["refid"], st["strand"], st["flag"], st["exon_off"], st["exon_st"], st["exon_en"], st["cigar_off"],
                           st["cigar"], st["mate"], st["group_off"], seq_off, seq_codes, st["src"], st["names"])
        out.append(b)

    cur = new_state()
    prev_name = None
    for ri, r in enumerate(recs):
        if r["flag"] & 4:
            continue
        idn = len(cur["refid"])
        new_name = (idn == 0) or (r["name"] != prev_name)
        if idn >= chunk_size and new_name:
            close(cur)
            cur = new_state()
            idn = 0
        if idn > 0 and new_name:
            cur["group_off"].append(idn)
        words = cigar_words(r["cigar"])
        ex = read_exons(r["pos"], words)
        cname = ann_chroms[r["chrom"]] if 0 <= r["chrom"] < len(ann_chroms) else None
        cur["refid"].append(index.refid_of.get(cname, -1))
        cur["strand"].append(ord("." if long_reads else read_strand(r, fr, rf)))
        cur["flag"].append(r["flag"])
        for s, e in ex:
            cur["exon_st"].append(s)
            cur["exon_en"].append(e)
        cur["exon_off"].append(len(cur["exon_st"]))
        cur["cigar"].extend(words)
        cur["cigar_off"].append(len(cur["cigar"]))
        cur["mate"].append(-1)
        cur["src"].append(ri)
        cur["names"].append(r["name"])
        cur["seqs"].append(r["seq"] or b"")
        start = r["pos"] + 1
        if (r["flag"] & 1) and r["chrom"] == r["mchrom"]:          # process_pairs
            mate_start = 0 if r["mpos"] < 0 else r["mpos"] + 1
            key = (r["name"], mate_start)
            h = cur["hashread"]
            if key in h:
                m = h.pop(key)
                if cur["mate"][idn] < 0:
                    cur["mate"][idn] = m
                if cur["mate"][m] < 0:
                    cur["mate"][m] = idn
            else:
                h[(r["name"], start)] = idn
        prev_name = r["name"]
    close(cur)
    return out
