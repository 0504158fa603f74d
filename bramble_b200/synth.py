"""Seeded synthetic inputs for the projection hot path (annotation, genome, alignments).

Shapes follow SURVEY.md section 8(d): GENCODE-like multi-isoform genes, paired-end short reads with a
small rate of clips / junction offsets / indels, ONT- or HiFi-like long reads with indel-rich CIGARs,
ragged exon boundaries, skipped or spurious small exons and soft clips whose bases come from the
neighbouring exon (so the ksw2 clip rescue of reference src/evaluate.cpp:451-656 has work to do).
Everything is produced from `numpy.random.default_rng(seed)`; no reference code is involved.

Records are plain dicts; `write_bam` serialises them (BGZF over zlib), `write_gtf` / `write_fasta`
write the annotation and the genome.  Used by tests/, bench.py and __graft_entry__.smoke().
"""
import struct
import zlib

import numpy as np

CIG = {c: i for i, c in enumerate("MIDNSHP=X")}
_NT16 = {c: i for i, c in enumerate("=ACMGRSVTWYHKDBN")}
_NT16_TABLE = bytes(_NT16.get(chr(c), 15) for c in range(256))      # letter -> 4-bit BAM code (anything else: N)


# ----------------------------------------------------------------------------- annotation
class Annotation:
    def __init__(self, chroms, chrom_len, tx):
        self.chroms = chroms          # names
        self.chrom_len = chrom_len    # lengths
        self.tx = tx                  # list of dict(id, gene, chrom, strand, exons=[(s,e)] 1-based incl.)

    def n_exons(self):
        return sum(len(t["exons"]) for t in self.tx)


def _lognorm(rng, median, sigma, lo, hi):
    return int(min(hi, max(lo, rng.lognormal(np.log(median), sigma))))


def make_annotation(seed, n_chrom=1, n_genes=500, iso_mean=4.0, iso_max=12, max_exons=20,
                    hi_iso_genes=0, hi_iso_exons=0, small_exon_rate=0.08, paralog_families=0,
                    chrom_prefix="chr"):
    """Genes laid out left to right on each chromosome; isoforms by exon skipping, alternative
    first/last exon ends and (rarely) alternative splice sites.  `hi_iso_genes` genes get 50-200
    isoforms and up to `hi_iso_exons` exons (forces the uint8 exon-index wrap of the reference,
    src/bramble.cpp:170).  `paralog_families` duplicates whole genes at other loci."""
    rng = np.random.default_rng(seed)
    chroms = [f"{chrom_prefix}{i + 1}" for i in range(n_chrom)]
    tx = []
    chrom_len = []
    per_chrom = [n_genes // n_chrom + (1 if i < n_genes % n_chrom else 0) for i in range(n_chrom)]
    gid = 0
    hi_left = hi_iso_genes
    par_left = paralog_families
    for ci in range(n_chrom):
        cursor = 10_000
        for _ in range(per_chrom[ci]):
            strand = "+" if rng.random() < 0.5 else "-"
            hi = hi_left > 0 and rng.random() < max(0.02, hi_iso_genes / max(1, n_genes))
            if hi:
                hi_left -= 1
            n_ex = int(rng.integers(1, max_exons + 1)) if not hi else int(rng.integers(max(2, hi_iso_exons // 2), hi_iso_exons + 1))
            exons = []
            pos = cursor
            for k in range(n_ex):
                if rng.random() < small_exon_rate and 0 < k < n_ex - 1:
                    ln = int(rng.integers(6, 36))
                else:
                    ln = _lognorm(rng, 150, 0.7, 30, 5000)
                exons.append((pos, pos + ln - 1))
                pos += ln + _lognorm(rng, 1500, 1.0, 60, 200_000)
            copies = 1
            if par_left > 0 and rng.random() < max(0.02, paralog_families / max(1, n_genes)):
                par_left -= 1
                copies = int(rng.integers(2, 8))
            n_iso = int(min(iso_max, 1 + rng.geometric(1.0 / iso_mean))) if not hi else int(rng.integers(50, 201))
            iso_list = []
            for ii in range(n_iso):
                if ii == 0 or n_ex <= 2:
                    ex = list(exons)
                else:
                    keep = [True] * n_ex
                    for k in range(1, n_ex - 1):
                        if rng.random() < 0.25:
                            keep[k] = False
                    lo = int(rng.integers(0, max(1, n_ex // 3)))
                    hi_k = n_ex - int(rng.integers(0, max(1, n_ex // 3)))
                    ex = [exons[k] for k in range(lo, hi_k) if keep[k]] or [exons[lo]]
                ex = [list(e) for e in ex]
                if ii > 0:
                    r = rng.random()
                    if r < 0.3:      # alternative 5'/3' end
                        ex[0][0] = max(cursor - 2000, ex[0][0] - int(rng.integers(1, 300)))
                    elif r < 0.5:
                        ex[-1][1] += int(rng.integers(1, 300))
                    elif r < 0.6 and len(ex) > 1:   # alternative splice site
                        k = int(rng.integers(0, len(ex) - 1))
                        d = int(rng.integers(1, 12))
                        if ex[k][1] - d > ex[k][0] + 5:
                            ex[k][1] -= d
                iso_list.append([tuple(e) for e in ex])
            # drop exact duplicate isoforms
            seen, uniq = set(), []
            for ex in iso_list:
                key = tuple(ex)
                if key not in seen:
                    seen.add(key)
                    uniq.append(ex)
            span = pos - cursor
            for cp in range(copies):
                shift = cp * (span + 5000)
                for ii, ex in enumerate(uniq):
                    tx.append(dict(id=f"G{gid}.c{cp}.t{ii}" if copies > 1 else f"G{gid}.t{ii}",
                                   gene=f"G{gid}" + (f".c{cp}" if copies > 1 else ""), chrom=ci, strand=strand,
                                   exons=[(s + shift, e + shift) for s, e in ex], family=gid))
            cursor = cursor + copies * (span + 5000) + _lognorm(rng, 8000, 1.0, 500, 500_000) - 5000 + 300
            cursor = max(cursor, max(e for t in tx[-len(uniq) * copies:] for _, e in t["exons"]) + 500)
            gid += 1
        chrom_len.append(cursor + 10_000)
    return Annotation(chroms, chrom_len, tx)


def write_gtf(ann, path, shuffle_seed=None):
    order = list(range(len(ann.tx)))
    if shuffle_seed is not None:
        np.random.default_rng(shuffle_seed).shuffle(order)
    with open(path, "w") as f:
        for i in order:
            t = ann.tx[i]
            c = ann.chroms[t["chrom"]]
            attr = f'gene_id "{t["gene"]}"; transcript_id "{t["id"]}";'
            f.write(f'{c}\tsynth\ttranscript\t{t["exons"][0][0]}\t{t["exons"][-1][1]}\t.\t{t["strand"]}\t.\t{attr}\n')
            for s, e in t["exons"]:
                f.write(f'{c}\tsynth\texon\t{s}\t{e}\t.\t{t["strand"]}\t.\t{attr}\n')


def make_genome(ann, seed, n_rate=0.0005):
    rng = np.random.default_rng(seed)
    out = []
    for L in ann.chrom_len:
        a = rng.integers(0, 4, size=L, dtype=np.uint8)
        s = np.frombuffer(b"ACGT", dtype=np.uint8)[a].copy()
        if n_rate > 0:
            s[rng.random(L) < n_rate] = ord("N")
        out.append(s.tobytes())
    return out


def write_fasta(ann, genome, path, width=60, with_fai=True):
    fai = []
    with open(path, "wb") as f:
        for name, seq in zip(ann.chroms, genome):
            f.write(b">" + name.encode() + b"\n")
            off = f.tell()
            for i in range(0, len(seq), width):
                f.write(seq[i:i + width] + b"\n")
            fai.append(f"{name}\t{len(seq)}\t{off}\t{width}\t{width + 1}\n")
    if with_fai:
        with open(path + ".fai", "w") as f:
            f.writelines(fai)


# ----------------------------------------------------------------------------- alignments
def _tx_len(t):
    return sum(e - s + 1 for s, e in t["exons"])


def _tx_segments(t, a, b):
    """Genomic segments (1-based inclusive, ascending) of [a,b) in GENOMIC-ASCENDING transcript
    coordinates (offset 0 = first base of the leftmost exon)."""
    segs, off = [], 0
    for s, e in t["exons"]:
        ln = e - s + 1
        lo, hi = max(a, off), min(b, off + ln)
        if lo < hi:
            segs.append([s + lo - off, s + hi - off - 1])
        off += ln
        if off >= b:
            break
    return segs


def _reg2bin(beg, end):
    end -= 1
    if beg >> 14 == end >> 14: return ((1 << 15) - 1) // 7 + (beg >> 14)
    if beg >> 17 == end >> 17: return ((1 << 12) - 1) // 7 + (beg >> 17)
    if beg >> 20 == end >> 20: return ((1 << 9) - 1) // 7 + (beg >> 20)
    if beg >> 23 == end >> 23: return ((1 << 6) - 1) // 7 + (beg >> 23)
    if beg >> 26 == end >> 26: return ((1 << 3) - 1) // 7 + (beg >> 26)
    return 0


def _segs_to_cigar(segs):
    cig = []
    for i, (s, e) in enumerate(segs):
        if i:
            cig.append(("N", s - segs[i - 1][1] - 1))
        cig.append(("M", e - s + 1))
    return cig


def _qlen(cig):
    return sum(l for o, l in cig if o in "MIS=X")


def _rlen(cig):
    return sum(l for o, l in cig if o in "MDN=X")


def _rand_seq(rng, n):
    return bytes(np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=n)])


def _norm_cigar(cig):
    out = []
    for o, l in cig:
        if l <= 0:
            continue
        if out and out[-1][0] == o:
            out[-1] = (o, out[-1][1] + l)
        else:
            out.append((o, l))
    return out


def _short_noise(rng, segs, p):
    """Returns cigar for a short read given its true segments; may perturb the segments."""
    segs = [list(s) for s in segs]
    lead, trail = [], []
    r = rng.random()
    if r < p["clip"]:                      # end soft clip (aligned part shrinks)
        k = int(rng.integers(1, 9))
        if rng.random() < 0.5 and segs[0][1] - segs[0][0] + 1 > k:
            segs[0][0] += k; lead = [("S", k)]
        elif segs[-1][1] - segs[-1][0] + 1 > k:
            segs[-1][1] -= k; trail = [("S", k)]
    elif r < p["clip"] + p["overhang"]:   # read runs past the exon boundary into intron/intergenic
        k = int(rng.integers(1, 9))
        if rng.random() < 0.5:
            segs[0][0] -= k
        else:
            segs[-1][1] += k
    elif r < p["clip"] + p["overhang"] + p["junc"] and len(segs) > 1:   # junction offset
        k = int(rng.integers(0, len(segs) - 1)); d = int(rng.integers(1, 4)) * (1 if rng.random() < 0.5 else -1)
        if rng.random() < 0.5:
            if segs[k][1] + d > segs[k][0]: segs[k][1] += d
        else:
            if segs[k + 1][0] + d < segs[k + 1][1]: segs[k + 1][0] += d
    cig = _segs_to_cigar(segs)
    r = rng.random()
    if r < p["indel"]:                     # small indel inside the longest M
        k = max(range(len(cig)), key=lambda i: cig[i][1] if cig[i][0] == "M" else -1)
        L = cig[k][1]
        if L > 12:
            at = int(rng.integers(4, L - 4)); n = int(rng.integers(1, 4))
            if rng.random() < 0.5:
                cig[k:k + 1] = [("M", at), ("I", n), ("M", L - at)]
            else:
                cig[k:k + 1] = [("M", at), ("D", n), ("M", L - at - n)] if L - at - n > 0 else [("M", L)]
    elif r < p["indel"] + p["eqx"]:
        cig = [("=" if o == "M" else o, l) for o, l in cig]
        k = [i for i, (o, l) in enumerate(cig) if o == "=" and l > 6]
        if k:
            i = k[0]; L = cig[i][1]; at = int(rng.integers(2, L - 2))
            cig[i:i + 1] = [("=", at), ("X", 1), ("=", L - at - 1)]
    if rng.random() < p["hard"]:
        lead = [("H", int(rng.integers(1, 30)))] + lead
    return segs, _norm_cigar(lead + cig + trail)


SHORT_DEFAULTS = dict(clip=0.02, overhang=0.015, junc=0.01, indel=0.01, eqx=0.01, hard=0.003,
                      intergenic=0.02, unspliced=0.02, unmapped=0.01, single=0.05, chimeric=0.02,
                      orphan=0.02, multi=0.08, multi_max=5, xs=0.9, anchor=0.05, diff_chrom=0.005)


def make_short_reads(ann, seed, n_names, read_len=150, stranded=None, params=None, multi_mean=None,
                     name_prefix="r"):
    """Name-collated paired-end short-read alignments.  stranded: None | 'fr' | 'rf' (orientation of
    read1/read2 made consistent with the reference's get_strand rule, src/bramble.cpp:213-244).
    Returns a list of record dicts in BAM order."""
    p = dict(SHORT_DEFAULTS); p.update(params or {})
    rng = np.random.default_rng(seed)
    tx = ann.tx
    lens = [_tx_len(t) for t in tx]
    by_family = {}
    for i, t in enumerate(tx):
        by_family.setdefault(t["family"], []).append(i)
    recs = []

    def one_end(t, a, b):
        segs = _tx_segments(t, a, b)
        return _short_noise(rng, segs, p)

    def mk(name, flag, chrom, segs, cig, mchrom, mpos, isize, xs):
        tags = [("NH", "C", 1), ("HI", "C", 1), ("AS", "C", int(rng.integers(100, 255))), ("NM", "C", int(rng.integers(0, 4)))]
        if xs:
            tags.insert(2, ("XS", "A", xs))
        ql = _qlen(cig)
        return dict(name=name, flag=flag, chrom=chrom, pos=segs[0][0] - 1, mapq=int(rng.choice([255, 3, 1, 0])),
                    cigar=cig, mchrom=mchrom, mpos=mpos, isize=isize, seq=_rand_seq(rng, ql),
                    qual=bytes(rng.integers(2, 41, size=ql, dtype=np.uint8)), tags=tags)

    for ni in range(n_names):
        name = f"{name_prefix}{ni:09d}"
        r = rng.random()
        if r < p["unmapped"]:
            recs.append(dict(name=name, flag=4, chrom=-1, pos=-1, mapq=0, cigar=[], mchrom=-1, mpos=-1, isize=0,
                             seq=_rand_seq(rng, read_len), qual=bytes([30] * read_len), tags=[]))
            continue
        n_aln = 1
        if multi_mean:
            n_aln = int(min(50, 1 + rng.geometric(1.0 / multi_mean) - 1)) or 1
        elif rng.random() < p["multi"]:
            n_aln = int(rng.integers(2, p["multi_max"] + 1))
        single = rng.random() < p["single"] if not multi_mean else rng.random() < 0.4
        ti0 = int(rng.integers(len(tx)))
        fam = by_family[tx[ti0]["family"]]
        for ai in range(n_aln):
            ti = ti0 if ai == 0 else (int(rng.choice(fam)) if rng.random() < 0.7 else int(rng.integers(len(tx))))
            t, L = tx[ti], lens[ti]
            sec = 256 if ai > 0 else 0
            rl = min(read_len, L)
            frag = int(min(L, max(rl, rng.normal(300, 50))))
            a = int(rng.integers(0, L - frag + 1))
            if rng.random() < p["anchor"]:    # start exactly on an exon boundary
                offs = np.cumsum([0] + [e - s + 1 for s, e in t["exons"]])[:-1]
                offs = [int(o) for o in offs if o + frag <= L]
                if offs:
                    a = int(rng.choice(offs))
            b = a + frag
            kind = rng.random()
            if kind < p["intergenic"]:       # alignment outside any exon
                s0 = int(rng.integers(1, 5000)); segsL = [[s0, s0 + rl - 1]]; cigL = [("M", rl)]
                segsR = [[s0 + 200, s0 + 199 + rl]]; cigR = [("M", rl)]
            elif kind < p["intergenic"] + p["unspliced"] and len(t["exons"]) > 1:   # retained intron
                k = int(rng.integers(0, len(t["exons"]) - 1)); s0 = max(1, t["exons"][k][1] - rl // 2)
                segsL = [[s0, s0 + rl - 1]]; cigL = [("M", rl)]
                segsR = [[s0 + 120, s0 + 119 + rl]]; cigR = [("M", rl)]
            else:
                segsL, cigL = one_end(t, a, a + rl)
                t2 = t
                if rng.random() < p["chimeric"]:
                    t2 = tx[int(rng.choice(fam))]
                    if _tx_len(t2) < b: t2 = t
                segsR, cigR = one_end(t2, b - rl, b)
            spl = len(segsL) > 1 or len(segsR) > 1
            xsL = t["strand"] if len(segsL) > 1 and rng.random() < p["xs"] else None
            xsR = t["strand"] if len(segsR) > 1 and rng.random() < p["xs"] else None
            chrom = t["chrom"]
            if single:
                rev = rng.random() < 0.5
                if stranded:   # unpaired: fr & !rev -> '-'
                    want_minus = t["strand"] == "-"
                    rev = (not want_minus) if stranded == "fr" else want_minus
                recs.append(mk(name, sec | (16 if rev else 0), chrom, segsL, cigL, -1, -1, 0, xsL))
                continue
            # left mate forward, right mate reverse.  Which is read1?
            if stranded == "fr":      # read1 forward => '-' transcript
                left_is_r1 = t["strand"] == "-"
            elif stranded == "rf":
                left_is_r1 = t["strand"] == "+"
            else:
                left_is_r1 = rng.random() < 0.5
            if stranded and rng.random() < 0.03:
                left_is_r1 = not left_is_r1
            fl = 1 | 2 | 32 | (64 if left_is_r1 else 128) | sec
            fr_ = 1 | 2 | 16 | (128 if left_is_r1 else 64) | sec
            posL, posR = segsL[0][0] - 1, segsR[0][0] - 1
            endR = segsR[-1][1]
            isz = endR - posL
            mchromR = chrom
            if rng.random() < p["diff_chrom"] and len(ann.chroms) > 1:
                mchromR = (chrom + 1) % len(ann.chroms)
            recL = mk(name, fl, chrom, segsL, cigL, mchromR, posR, isz, xsL)
            recR = mk(name, fr_, mchromR, segsR, cigR, chrom, posL, -isz, xsR)
            if rng.random() < p["orphan"]:     # mate missing from the file
                recs.append(recL if rng.random() < 0.5 else recR)
            elif rng.random() < 0.5:
                recs.extend([recL, recR])
            else:
                recs.extend([recR, recL])
    return recs


LONG_DEFAULTS = dict(err=0.08, clip=0.5, clip_true=0.7, ragged=0.5, ragged_max=48, junc_shift=0.15,
                     junc_max=14, skip_small=0.5, spurious=0.04, multi=0.1, unmapped=0.01, partial=0.4,
                     intergenic=0.02, no_seq_secondary=0.3, hard=0.05)


def _tx_bases(t, genome):
    g = genome[t["chrom"]]
    return b"".join(g[s - 1:e] for s, e in t["exons"])


def make_long_reads(ann, genome, seed, n_names, params=None, median_len=1200, name_prefix="L"):
    """Unpaired long-read alignments (minimap2-like: ts:A tag, 0/16/256/2048 flags).  Soft-clipped
    bases are, with probability clip_true, the true transcript bases adjacent to the aligned part
    (a few errors added), so that the clip rescue can extend into the neighbouring exon."""
    p = dict(LONG_DEFAULTS); p.update(params or {})
    rng = np.random.default_rng(seed)
    tx = ann.tx
    lens = [_tx_len(t) for t in tx]
    by_family = {}
    for i, t in enumerate(tx):
        by_family.setdefault(t["family"], []).append(i)
    recs = []
    comp = bytes.maketrans(b"ACGTN", b"TGCAN")

    def errorful(cig_len):
        """Split an aligned block of cig_len reference bases into M/I/D/X runs."""
        out, i = [], 0
        rate = p["err"]
        if rate <= 0 or cig_len < 8:
            return [("M", cig_len)]
        while i < cig_len:
            gap = int(rng.geometric(rate)) if rate > 0 else cig_len
            run = min(cig_len - i, gap)
            out.append(("M", run)); i += run
            if i >= cig_len:
                break
            r = rng.random(); n = int(rng.integers(1, 4))
            if r < 0.35:
                out.append(("I", n))
            elif r < 0.7:
                n = min(n, cig_len - i)
                if n > 0 and cig_len - i - n > 0:
                    out.append(("D", n)); i += n
            else:
                out.append(("X", 1)); i += 1
        # blocks must start and end on a match
        while out and out[0][0] != "M": out.pop(0)
        while out and out[-1][0] != "M": out.pop()
        need = cig_len - _rlen(out)
        if need > 0: out.append(("M", need))
        return _norm_cigar([("M" if o == "X" and rng.random() < 0.5 else o, l) for o, l in out])

    for ni in range(n_names):
        name = f"{name_prefix}{ni:09d}"
        if rng.random() < p["unmapped"]:
            recs.append(dict(name=name, flag=4, chrom=-1, pos=-1, mapq=0, cigar=[], mchrom=-1, mpos=-1, isize=0,
                             seq=_rand_seq(rng, 200), qual=b"\xff" * 200, tags=[]))
            continue
        n_aln = int(rng.integers(2, 5)) if rng.random() < p["multi"] else 1
        ti0 = int(rng.integers(len(tx)))
        fam = by_family[tx[ti0]["family"]]
        group = []
        for ai in range(n_aln):
            ti = ti0 if ai == 0 else int(rng.choice(fam))
            t, L = tx[ti], lens[ti]
            nex = len(t["exons"])
            offs = np.cumsum([0] + [e - s + 1 for s, e in t["exons"]])
            if rng.random() < p["intergenic"]:
                s0 = int(rng.integers(1, 5000)); segs = [[s0, s0 + 499]]
                a, b = 0, 500
                ek0 = ek1 = 0
            else:
                want = int(min(L, max(60, rng.lognormal(np.log(median_len), 0.6))))
                if rng.random() < p["partial"] or want >= L:
                    a, b = 0, L
                else:
                    a = int(rng.integers(0, L - want + 1)); b = a + want
                # exon-anchored ends help the clip rescue: snap to exon starts/ends half of the time
                if rng.random() < 0.6:
                    k = int(np.searchsorted(offs, a, side="right") - 1); a = int(offs[k])
                if rng.random() < 0.6:
                    k = int(np.searchsorted(offs, b - 1, side="right")); b = int(offs[k])
                if b - a < 30: a, b = 0, L
                segs = _tx_segments(t, a, b)
                ek0 = int(np.searchsorted(offs, a, side="right") - 1)
            true_left = _tx_bases(t, genome)[:a] if genome else b""
            true_right = _tx_bases(t, genome)[b:] if genome else b""
            # skipped small exon (guide exon missing from the read)
            if len(segs) > 2 and rng.random() < p["skip_small"]:
                small = [k for k in range(1, len(segs) - 1) if segs[k][1] - segs[k][0] + 1 <= 35]
                if small:
                    del segs[int(rng.choice(small))]
            # spurious small exon inside an intron
            if len(segs) > 1 and rng.random() < p["spurious"]:
                k = int(rng.integers(0, len(segs) - 1))
                gap = segs[k + 1][0] - segs[k][1] - 1
                if gap > 200:
                    ln = int(rng.integers(5, 45)); s0 = segs[k][1] + int(rng.integers(50, gap - ln - 50))
                    segs.insert(k + 1, [s0, s0 + ln - 1])
            # junction shifts
            for k in range(len(segs) - 1):
                if rng.random() < p["junc_shift"]:
                    d = int(rng.integers(-p["junc_max"], p["junc_max"] + 1))
                    if rng.random() < 0.5:
                        if segs[k][1] + d > segs[k][0] + 3 and segs[k][1] + d < segs[k + 1][0] - 3: segs[k][1] += d
                    else:
                        if segs[k + 1][0] + d < segs[k + 1][1] - 3 and segs[k + 1][0] + d > segs[k][1] + 3: segs[k + 1][0] += d
            lead, trail = [], []
            clipL = clipR = 0
            if rng.random() < p["clip"]:
                clipL = int(min(400, rng.geometric(1 / 30.0)))
            if rng.random() < p["clip"]:
                clipR = int(min(400, rng.geometric(1 / 30.0)))
            # ragged outer ends (only when not clipped there, keeps rescue anchors exact)
            if not clipL and rng.random() < p["ragged"]:
                d = int(rng.integers(-p["ragged_max"], p["ragged_max"] + 1))
                if segs[0][0] + d < segs[0][1] - 3 and segs[0][0] + d > 0: segs[0][0] += d
            if not clipR and rng.random() < p["ragged"]:
                d = int(rng.integers(-p["ragged_max"], p["ragged_max"] + 1))
                if segs[-1][1] + d > segs[-1][0] + 3: segs[-1][1] += d
            cig = []
            for k, (s, e) in enumerate(segs):
                if k:
                    cig.append(("N", s - segs[k - 1][1] - 1))
                cig.extend(errorful(e - s + 1))
            if clipL: lead = [("S", clipL)]
            if clipR: trail = [("S", clipR)]
            hard = rng.random() < p["hard"]
            if hard:
                lead = [("H", int(rng.integers(5, 60)))] + lead
                if rng.random() < 0.5: trail = trail + [("H", int(rng.integers(5, 60)))]
            cig = _norm_cigar(lead + cig + trail)
            # sequence (genomic forward orientation)
            if genome:
                g = genome[t["chrom"]]
                parts, gp = [], segs[0][0]
                for o, l in cig:
                    if o in "M=":
                        parts.append(g[gp - 1:gp - 1 + l]); gp += l
                    elif o == "X":
                        parts.append(_rand_seq(rng, l)); gp += l
                    elif o in "DN":
                        gp += l
                    elif o == "I":
                        parts.append(_rand_seq(rng, l))
                    elif o == "S":
                        left = not parts
                        if rng.random() < p["clip_true"]:
                            src = true_left[-l:] if left else true_right[:l]
                            src = bytearray(src)
                            for x in range(len(src)):
                                if rng.random() < p["err"] / 2: src[x] = b"ACGT"[int(rng.integers(4))]
                            pad = _rand_seq(rng, l - len(src))
                            parts.append((pad + bytes(src)) if left else (bytes(src) + pad))
                        else:
                            parts.append(_rand_seq(rng, l))
                seq = b"".join(parts)
            else:
                seq = _rand_seq(rng, _qlen(cig))
            rev = rng.random() < 0.5
            flag = (16 if rev else 0) | (256 if ai > 0 and rng.random() < 0.7 else (2048 if ai > 0 else 0))
            qual = bytes(rng.integers(2, 41, size=len(seq), dtype=np.uint8))
            if (flag & 256) and rng.random() < p["no_seq_secondary"]:
                seq, qual = b"", b""      # minimap2 writes '*' for secondaries
            tags = [("NM", "S" if len(seq) > 255 else "C", int(rng.integers(0, 200))), ("AS", "i", int(rng.integers(50, 3000))),
                    ("ts", "A", "+" if rng.random() < 0.5 else "-"), ("NH", "C", n_aln), ("HI", "C", ai + 1)]
            group.append(dict(name=name, flag=flag, chrom=t["chrom"], pos=segs[0][0] - 1, mapq=int(rng.integers(0, 61)),
                              cigar=cig, mchrom=-1, mpos=-1, isize=0, seq=seq, qual=qual, tags=tags))
        recs.extend(group)
    return recs


# ----------------------------------------------------------------------------- BAM writer
def _pack_record(r, compute_bin=True):
    name = r["name"].encode() + b"\0"
    cig = r["cigar"]
    ncig = len(cig)
    cigb = struct.pack(f"<{ncig}I", *[(l << 4) | CIG[o] for o, l in cig]) if ncig else b""
    seq = r["seq"] or b""
    lq = len(seq)
    if lq:
        codes = np.frombuffer(bytes(seq).translate(_NT16_TABLE) + (b"\0" if lq & 1 else b""), dtype=np.uint8)
        seqb = ((codes[0::2] << 4) | codes[1::2]).astype(np.uint8).tobytes()
    else:
        seqb = b""
    qual = r["qual"] if r.get("qual") else b"\xff" * lq
    aux = bytearray()
    for tag, ty, val in r.get("tags", []):
        aux += tag.encode() + ty.encode()
        if ty == "A": aux += val.encode()
        elif ty in ("Z", "H"): aux += val.encode() + b"\0"
        elif ty == "B":                                  # (subtype, values)
            sub, vals = val
            aux += sub.encode() + struct.pack("<I", len(vals)) + struct.pack("<%d%s" % (len(vals), {"c": "b", "C": "B", "s": "h", "S": "H", "i": "i", "I": "I", "f": "f"}[sub]), *vals)
        else: aux += struct.pack("<" + {"c": "b", "C": "B", "s": "h", "S": "H", "i": "i", "I": "I", "f": "f"}[ty], val)
    rl = _rlen(cig) or 1
    b_ = _reg2bin(r["pos"], r["pos"] + rl) if (compute_bin and r["pos"] >= 0) else 4680
    core = struct.pack("<iiIIiiii", r["chrom"], r["pos"], (b_ << 16) | (r["mapq"] << 8) | len(name),
                       (r["flag"] << 16) | ncig, lq, r["mchrom"], r["mpos"], r["isize"])
    body = core + name + cigb + seqb + qual + bytes(aux)
    return struct.pack("<i", len(body)) + body


def _bgzf_block(data, level):
    co = zlib.compressobj(level, zlib.DEFLATED, -15)
    c = co.compress(data) + co.flush()
    bs = len(c) + 25
    return (b"\x1f\x8b\x08\x04\0\0\0\0\0\xff\x06\0BC\x02\0" + struct.pack("<H", bs) + c +
            struct.pack("<II", zlib.crc32(data) & 0xffffffff, len(data)))


_BGZF_EOF = bytes([0x1f, 0x8b, 8, 4, 0, 0, 0, 0, 0, 0xff, 6, 0, 0x42, 0x43, 2, 0, 0x1b, 0, 3, 0, 0, 0, 0, 0, 0, 0, 0, 0])


def write_bam(path, ann, recs, extra_header="@PG\tID:synth\tPN:synth\tVN:1\n@CO\tsynthetic alignments\n", level=1,
              hd="@HD\tVN:1.6\tSO:unsorted\tGO:query\n"):
    text = hd + "".join(f"@SQ\tSN:{n}\tLN:{l}\n" for n, l in zip(ann.chroms, ann.chrom_len)) + extra_header
    tb = text.encode()
    head = bytearray(b"BAM\1" + struct.pack("<i", len(tb)) + tb + struct.pack("<i", len(ann.chroms)))
    for n, l in zip(ann.chroms, ann.chrom_len):
        nb = n.encode() + b"\0"
        head += struct.pack("<i", len(nb)) + nb + struct.pack("<i", l)
    with open(path, "wb") as f:
        buf = bytearray(head)
        for r in recs:
            rec = _pack_record(r)
            if len(buf) + len(rec) > 0xff00 and buf:
                f.write(_bgzf_block(bytes(buf), level)); buf = bytearray()
            buf += rec
            while len(buf) > 0xff00:
                f.write(_bgzf_block(bytes(buf[:0xff00]), level)); del buf[:0xff00]
        if buf:
            f.write(_bgzf_block(bytes(buf), level))
        f.write(_BGZF_EOF)


def read_bam(path):
    """Decode a BAM file into (header_text, refs, [record dict with raw fields]).  Test helper used
    to diff product output against oracle output record by record."""
    raw = bytearray()
    with open(path, "rb") as f:
        data = f.read()
    p = 0
    while p < len(data):
        xlen = struct.unpack_from("<H", data, p + 10)[0]
        bsize = struct.unpack_from("<H", data, p + 16)[0]   # files written here/by the shims: BC is the first subfield
        cstart = p + 12 + xlen
        cend = p + bsize + 1 - 8
        raw += zlib.decompress(data[cstart:cend], -15)
        p += bsize + 1
    assert raw[:4] == b"BAM\1"
    lt = struct.unpack_from("<i", raw, 4)[0]
    text = raw[8:8 + lt].decode()
    q = 8 + lt
    nref = struct.unpack_from("<i", raw, q)[0]; q += 4
    refs = []
    for _ in range(nref):
        ln = struct.unpack_from("<i", raw, q)[0]; q += 4
        nm = raw[q:q + ln - 1].decode(); q += ln
        refs.append((nm, struct.unpack_from("<i", raw, q)[0])); q += 4
    recs = []
    while q < len(raw):
        bl = struct.unpack_from("<i", raw, q)[0]
        recs.append(bytes(raw[q + 4:q + 4 + bl])); q += 4 + bl
    return text, refs, recs


def describe_record(rec, refs=None):
    """Human-readable one-line SAM-ish view of a raw BAM record body (debugging / diffs)."""
    tid, pos, bmq, fnc, lq, mtid, mpos, isz = struct.unpack_from("<iiIIiiii", rec, 0)
    lname = bmq & 0xff; ncig = fnc & 0xffff
    name = rec[32:32 + lname - 1].decode()
    cig = struct.unpack_from(f"<{ncig}I", rec, 32 + lname)
    cs = "".join(f"{c >> 4}{'MIDNSHP=XB??????'[c & 15]}" for c in cig)
    aux = rec[32 + lname + 4 * ncig + (lq + 1) // 2 + lq:]
    return f"{name}\t{fnc >> 16}\t{tid}\t{pos}\t{(bmq >> 8) & 0xff}\t{cs}\t{mtid}\t{mpos}\t{isz}\tlq={lq}\tbin={bmq >> 16}\taux={aux.hex()}"
