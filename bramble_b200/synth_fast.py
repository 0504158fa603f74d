"""Vectorised (numpy) synthetic workloads at benchmark scale: a GENCODE-shaped annotation and
name-collated paired-end short reads, produced directly as the engine's structure-of-arrays
(`brb_index_desc` / `brb_batch_in`) without going through BAM.  Same shapes as SURVEY.md 8(d) C1/C2/C5;
`to_records()` turns a slice back into BAM record dicts (for the reference arm of bench.py) and
`write_gtf()` writes the matching annotation.  Seeded; no reference code involved."""
import numpy as np

from . import capi, pack


class FastAnnotation:
    pass


def _seg_cumsum_excl(vals, seg_start_idx, seg_id):
    """exclusive cumulative sum restarting at every segment"""
    c = np.cumsum(vals)
    base = (c - vals)[seg_start_idx]
    return c - vals - base[seg_id]


def make_annotation(seed, n_chrom=25, n_genes=60000, iso_mean=4.2, iso_cap=200, max_exons=20, hi_iso_genes=0, hi_iso_exons=300,
                    paralog_families=0, paralog_cap=50):
    """GENCODE-shaped annotation.  Gene TEMPLATES (exon lengths, isoforms by exon skipping, alternative ends) are drawn
    first; every template is laid out at one locus, or -- `paralog_families` of them -- at 2..paralog_cap loci with fresh
    introns (transcripts duplicated at 2-50 loci).  `hi_iso_genes` templates get 50-200 isoforms and
    hi_iso_exons/2..hi_iso_exons exons (C4: forces the reference's uint8 exon_id wrap, src/bramble.cpp:170).  The optional
    extras draw from their own generator, so the plain shape of a seed never changes."""
    rng = np.random.default_rng(seed)
    rng2 = np.random.default_rng(seed + 7919)
    A = FastAnnotation()
    A.chroms = [f"chr{i + 1}" for i in range(n_chrom)]
    g_chrom = np.sort(rng.integers(0, n_chrom, size=n_genes)).astype(np.int64)
    g_nex = rng.integers(1, max_exons + 1, size=n_genes).astype(np.int64)
    g_strand = rng.random(n_genes) < 0.5                      # True = '-'
    hi = None
    if hi_iso_genes:
        hi = rng2.choice(n_genes, size=min(hi_iso_genes, n_genes), replace=False)
        g_nex[hi] = rng2.integers(max(2, hi_iso_exons // 2), hi_iso_exons + 1, size=hi.size)
    g_eoff = np.concatenate([[0], np.cumsum(g_nex)])
    E = int(g_eoff[-1])
    e_gene = np.repeat(np.arange(n_genes), g_nex)
    e_k = np.arange(E) - g_eoff[e_gene]
    e_len = np.clip(rng.lognormal(np.log(150), 0.7, size=E), 30, 5000).astype(np.int64)
    small = (rng.random(E) < 0.08) & (e_k > 0) & (e_k < g_nex[e_gene] - 1)
    e_len[small] = rng.integers(6, 36, size=int(small.sum()))
    intron = np.clip(rng.lognormal(np.log(1500), 1.0, size=E), 60, 200000).astype(np.int64)
    inter = np.clip(rng.lognormal(np.log(8000), 1.0, size=n_genes), 1500, 500000).astype(np.int64)
    # isoforms of the templates: (template transcript, exon) keep mask and alternative transcript ends
    g_niso = np.minimum(iso_cap, rng.geometric(1.0 / iso_mean, size=n_genes)).astype(np.int64)
    if hi is not None:
        g_niso[hi] = rng2.integers(50, 201, size=hi.size)
    g_toff = np.concatenate([[0], np.cumsum(g_niso)])
    TT = int(g_toff[-1])
    tt_gene = np.repeat(np.arange(n_genes), g_niso)
    tt_iso = np.arange(TT) - g_toff[tt_gene]
    tt_nge = g_nex[tt_gene]
    pt_off = np.concatenate([[0], np.cumsum(tt_nge)])
    PT = int(pt_off[-1])
    pt_t = np.repeat(np.arange(TT), tt_nge)
    pt_k = np.arange(PT) - pt_off[pt_t]
    edge = (pt_k == 0) | (pt_k == tt_nge[pt_t] - 1)
    keep_tpl = edge | (tt_iso[pt_t] == 0) | (rng.random(PT) >= 0.25)
    r = rng.random(TT)
    alt5 = (tt_iso > 0) & (r < 0.3)
    alt3 = (tt_iso > 0) & (r >= 0.3) & (r < 0.5)
    d5 = np.zeros(TT, dtype=np.int64); d3 = np.zeros(TT, dtype=np.int64)
    d5[alt5] = rng.integers(1, 300, size=int(alt5.sum()))
    d3[alt3] = rng.integers(1, 300, size=int(alt3.sum()))
    # gene instances: one per template, paralog families at several loci
    copies = np.ones(n_genes, dtype=np.int64)
    if paralog_families:
        fam = rng2.choice(n_genes, size=min(paralog_families, n_genes), replace=False)
        copies[fam] = np.clip(1 + rng2.geometric(1.0 / 4.0, size=fam.size), 2, paralog_cap)
    i_tpl = np.repeat(np.arange(n_genes), copies)
    i_copy = np.arange(i_tpl.size) - (np.cumsum(copies) - copies)[i_tpl]
    i_chrom = g_chrom[i_tpl].copy()
    extra = i_copy > 0
    i_chrom[extra] = rng2.integers(0, n_chrom, size=int(extra.sum()))
    order = np.argsort(i_chrom, kind="stable")
    i_tpl, i_copy, i_chrom, extra = i_tpl[order], i_copy[order], i_chrom[order], extra[order]
    n_inst = int(i_tpl.size)
    i_nex = g_nex[i_tpl]
    i_eoff = np.concatenate([[0], np.cumsum(i_nex)])
    EI = int(i_eoff[-1])
    ie_inst = np.repeat(np.arange(n_inst), i_nex)
    ie_k = np.arange(EI) - i_eoff[ie_inst]
    ie_src = g_eoff[i_tpl[ie_inst]] + ie_k
    ie_len = e_len[ie_src]
    ie_intron = intron[ie_src].copy()
    i_inter = inter[i_tpl].copy()
    if extra.any():
        xe_ = extra[ie_inst]
        ie_intron[xe_] = np.clip(rng2.lognormal(np.log(1500), 1.0, size=int(xe_.sum())), 60, 200000).astype(np.int64)
        i_inter[extra] = np.clip(rng2.lognormal(np.log(8000), 1.0, size=int(extra.sum())), 1500, 500000).astype(np.int64)
    # gap BEFORE each exon: intergenic for the first exon of a gene, intron otherwise
    gap = np.where(ie_k == 0, i_inter[ie_inst], ie_intron)
    step = gap + ie_len
    # positions restart on every chromosome
    e_chrom = i_chrom[ie_inst]
    first_of_chrom = np.concatenate([[True], e_chrom[1:] != e_chrom[:-1]])
    cidx = np.nonzero(first_of_chrom)[0]
    cid = np.cumsum(first_of_chrom) - 1
    endpos = _seg_cumsum_excl(step, cidx, cid) + step + 10000       # 1-based inclusive end
    e_s = endpos - ie_len + 1
    e_e = endpos
    cmax = np.zeros(n_chrom, dtype=np.int64)
    np.maximum.at(cmax, e_chrom, e_e)
    A.chrom_len = [int(m) + 10000 for m in cmax]
    # transcripts of the instances
    it_niso = g_niso[i_tpl]
    T = int(it_niso.sum())
    t_gene = np.repeat(np.arange(n_inst), it_niso)
    t_iso = np.arange(T) - np.concatenate([[0], np.cumsum(it_niso)])[t_gene]
    t_tpl = g_toff[i_tpl[t_gene]] + t_iso                      # template transcript
    # (transcript, gene exon) pairs
    t_nge = i_nex[t_gene]
    p_off = np.concatenate([[0], np.cumsum(t_nge)])
    P = int(p_off[-1])
    p_t = np.repeat(np.arange(T), t_nge)
    p_k = np.arange(P) - p_off[p_t]
    p_e = i_eoff[t_gene[p_t]] + p_k
    keep = keep_tpl[pt_off[t_tpl[p_t]] + p_k]
    x_t = p_t[keep]
    x_s = e_s[p_e[keep]].copy()
    x_e = e_e[p_e[keep]].copy()
    t_nex = np.bincount(x_t, minlength=T).astype(np.int64)
    x_off = np.concatenate([[0], np.cumsum(t_nex)])
    # alternative transcript ends on non-first isoforms (outward, stays inside the intergenic gap)
    x_s[x_off[:-1]] -= d5[t_tpl]
    x_e[x_off[1:] - 1] += d3[t_tpl]
    A.n_tx = T
    A.t_gene, A.t_iso = t_gene, t_iso
    A.g_family, A.g_copies = i_tpl, copies[i_tpl]              # per gene instance: its template and how many loci the template has
    fam_order = np.argsort(i_tpl, kind="stable")               # instances of template f: fam_inst[fam_off[f] : fam_off[f + 1]]
    A.fam_inst, A.fam_off = fam_order, np.concatenate([[0], np.cumsum(copies)])
    A.t_chrom = i_chrom[t_gene]
    A.t_minus = g_strand[i_tpl[t_gene]]
    A.t_nex, A.x_off, A.x_s, A.x_e = t_nex, x_off, x_s, x_e          # exons genomic-ascending, 1-based inclusive
    A.t_start = x_s[x_off[:-1]]
    A.t_end = x_e[x_off[1:] - 1]
    x_len = x_e - x_s + 1
    A.x_len = x_len
    A.x_cum = _seg_cumsum_excl(x_len, x_off[:-1], np.repeat(np.arange(T), t_nex))   # offset of exon in genomic-ascending tx coords
    A.t_len = np.add.reduceat(x_len, x_off[:-1])
    A.ids = None
    return A


def tx_ids(A):
    if A.ids is None:
        A.ids = [f"G{g:07d}.t{i:03d}" for g, i in zip(A.t_gene.tolist(), A.t_iso.tolist())]
    return A.ids


def tid_order(A):
    """gclib gfo_cmpByLoc (gff.cpp:75-91): chromosome NAME (strcmp), start, level, end, ID."""
    crank = np.argsort(np.argsort(np.array(A.chroms)))      # lexicographic rank of chromosome names
    # IDs are zero padded, so (gene, iso) numeric order == strcmp order
    return np.lexsort((A.t_iso, A.t_gene, A.t_end, A.t_start, crank[A.t_chrom]))


def to_annotation(A):
    """The same annotation as a synth.Annotation (list of transcript dicts), for the per-record generators."""
    from . import synth
    ids = tx_ids(A)
    xs, xe, xo = A.x_s.tolist(), A.x_e.tolist(), A.x_off.tolist()
    tg, tc, tm = A.t_gene.tolist(), A.t_chrom.tolist(), A.t_minus.tolist()
    tx = [dict(id=ids[t], gene=f"G{tg[t]:07d}", chrom=tc[t], strand="-" if tm[t] else "+",
               exons=list(zip(xs[xo[t]:xo[t + 1]], xe[xo[t]:xo[t + 1]])), family=tg[t]) for t in range(A.n_tx)]
    return synth.Annotation(list(A.chroms), list(A.chrom_len), tx)


def make_genome_codes(A, seed, n_rate=0.0005):
    """Random genome as ksw2 base codes (0..3 = ACGT, 4 = N), one uint8 array per chromosome."""
    rng = np.random.default_rng(seed)
    out = []
    for L in A.chrom_len:
        g = rng.integers(0, 4, size=L, dtype=np.uint8)
        if n_rate > 0:
            g[rng.random(L) < n_rate] = 4
        out.append(g)
    return out


def genome_bytes(codes):
    lut = np.frombuffer(b"ACGTN", dtype=np.uint8)
    return [lut[c].tobytes() for c in codes]


def write_fasta(A, codes, path, width=60):
    lut = np.frombuffer(b"ACGTN", dtype=np.uint8)
    fai = []
    with open(path, "wb") as f:
        for name, c in zip(A.chroms, codes):
            f.write(b">" + name.encode() + b"\n")
            off = f.tell()
            s = lut[c]
            full = (len(s) // width) * width
            body = np.empty((full // width, width + 1), dtype=np.uint8)
            body[:, :width] = s[:full].reshape(-1, width)
            body[:, width] = 10
            f.write(body.tobytes())
            if full < len(s):
                f.write(s[full:].tobytes() + b"\n")
            fai.append(f"{name}\t{len(s)}\t{off}\t{width}\t{width + 1}\n")
    with open(path + ".fai", "w") as f:
        f.writelines(fai)


def build_index(A, genome_codes=None):
    """Vectorised twin of pack.build_index (reference src/bramble.cpp:132-211).  genome_codes: per-chromosome uint8 codes
    (make_genome_codes), indexed like A.chroms."""
    ix = pack.Index()
    order = tid_order(A)                       # tid -> transcript
    T = A.n_tx
    tid_of = np.empty(T, dtype=np.int64); tid_of[order] = np.arange(T)
    ids = tx_ids(A)
    ix.tid_names = [ids[i] for i in order.tolist()]
    ix.tid_len = A.t_len[order].tolist()
    ix.tid_strand = ["-" if m else "+" for m in A.t_minus[order].tolist()]
    # refid = first appearance in the GTF, which write_gtf emits in transcript (generation) order
    seen = []
    for c in A.t_chrom[np.sort(np.unique(A.t_chrom, return_index=True)[1])].tolist():
        seen.append(c)
    for c in seen:
        ix.refid_of[A.chroms[c]] = len(ix.refid_of)
    refid_of_chrom = np.full(len(A.chroms), -1, dtype=np.int64)
    for c in seen:
        refid_of_chrom[c] = ix.refid_of[A.chroms[c]]
    n_ref = len(seen)
    # exons in tid order, transcript (5'->3') order inside a transcript
    nex = A.t_nex[order]
    off = np.concatenate([[0], np.cumsum(nex)])
    X = int(off[-1])
    r_t = np.repeat(np.arange(T), nex)                         # tid of each row
    r_k = np.arange(X) - off[r_t]                              # position in transcript order
    src_t = order[r_t]
    minus = A.t_minus[src_t]
    gidx = np.where(minus, nex[r_t] - 1 - r_k, r_k)            # genomic exon index
    src_x = A.x_off[src_t] + gidx
    st = A.x_s[src_x]; en = A.x_e[src_x] + 1
    ln = en - st
    pos_start = _seg_cumsum_excl(ln, off[:-1], r_t)
    tlen = A.t_len[src_t]
    has_prev = r_k > 0; has_next = r_k < nex[r_t] - 1
    prev_st = np.where(has_prev, np.roll(st, 1), 0); prev_en = np.where(has_prev, np.roll(en, 1), 0)
    next_st = np.where(has_next, np.roll(st, -1), 0); next_en = np.where(has_next, np.roll(en, -1), 0)
    table = 2 * refid_of_chrom[A.t_chrom[src_t]] + minus
    # group rows by table keeping insertion order (stable), then the reference's std::sort by start per table
    ordr = np.argsort(table, kind="stable")
    counts = np.bincount(table, minlength=2 * n_ref)
    table_off = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint64)
    st_g = st[ordr].astype(np.int32)
    perm = np.empty(X, dtype=np.int64)
    for t in range(2 * n_ref):
        lo, hi = int(table_off[t]), int(table_off[t + 1])
        if hi > lo:
            perm[lo:hi] = lo + pack.sort_perm_by_start(np.ascontiguousarray(st_g[lo:hi])).astype(np.int64)
    fin = ordr[perm]
    cols = dict(st=st[fin].astype(np.int32), en=en[fin].astype(np.int32), tid=r_t[fin].astype(np.uint32),
                exon_id=(gidx[fin] & 0xff).astype(np.uint8), pos_start=pos_start[fin].astype(np.uint32),
                prev_st=prev_st[fin].astype(np.uint32), prev_en=prev_en[fin].astype(np.uint32),
                next_st=next_st[fin].astype(np.uint32), next_en=next_en[fin].astype(np.uint32),
                has_prev_next=(has_prev[fin].astype(np.uint8) | (has_next[fin].astype(np.uint8) << 1)),
                transcript_len=tlen[fin].astype(np.uint32))
    d = ix.desc
    ix.arrays = {k: np.ascontiguousarray(v) for k, v in cols.items()}
    ix.arrays["table_off"] = table_off
    d.n_tables = 2 * n_ref
    d.table_off = capi.ptr(ix.arrays["table_off"])
    for k in cols:
        setattr(d, k, capi.ptr(ix.arrays[k]))
    d.n_chrom = 0; d.chrom_off = None; d.genome_codes = None
    if genome_codes is not None:
        by_ref = [np.zeros(0, np.uint8)] * n_ref
        for c in seen:
            by_ref[ix.refid_of[A.chroms[c]]] = genome_codes[c]
        flat = np.concatenate(by_ref + [np.full(1, 4, np.uint8)])                    # (pad to an even count)
        nb = int(sum(len(g) for g in by_ref))
        flat = flat[:nb + (nb & 1)]
        ix.arrays["genome_codes"] = np.ascontiguousarray(flat[0::2] | (flat[1::2] << 4))   # two codes per byte: brb_index_desc::genome_format 1
        ix.arrays["chrom_off"] = np.concatenate([[0], np.cumsum([len(g) for g in by_ref])]).astype(np.uint64)
        d.n_chrom = n_ref; d.genome_format = 1
        d.chrom_off = capi.ptr(ix.arrays["chrom_off"]); d.genome_codes = capi.ptr(ix.arrays["genome_codes"])
    ix.refid_of_chrom = refid_of_chrom
    return ix


def write_gtf(A, path):
    ids = tx_ids(A)
    with open(path, "w") as f:
        xs, xe, xo = A.x_s.tolist(), A.x_e.tolist(), A.x_off.tolist()
        for t in range(A.n_tx):
            c = A.chroms[int(A.t_chrom[t])]; sd = "-" if A.t_minus[t] else "+"
            attr = f'gene_id "G{int(A.t_gene[t]):07d}"; transcript_id "{ids[t]}";'
            lines = [f"{c}\tsynth\ttranscript\t{xs[xo[t]]}\t{xe[xo[t + 1] - 1]}\t.\t{sd}\t.\t{attr}\n"]
            for k in range(xo[t], xo[t + 1]):
                lines.append(f"{c}\tsynth\texon\t{xs[k]}\t{xe[k]}\t.\t{sd}\t.\t{attr}\n")
            f.write("".join(lines))


# ----------------------------------------------------------------------------- reads
class FastBatch(pack.Batch):
    """pack.Batch plus the columns needed to turn alignments back into BAM records."""
    pass


def make_short_batch(A, ix, seed, n_names, read_len=150, stranded="fr", multi=0.08, multi_mean=None, noise=True, name_base=0,
                     paired_frac=1.0):
    """Name-collated paired-end short reads as one SoA bundle.  stranded: 'fr' | 'rf' | None.
    multi: fraction of names with extra (secondary) fragments; multi_mean: if set, every name has
    1+Geometric(multi_mean) fragments capped at 50 (the C5 multi-mapping stress shape).  On an annotation with paralog
    families a secondary fragment of a family gene lands on the same isoform at another locus of the family (a true
    multi-mapper).  paired_frac < 1: the remaining names are single-end (one alignment per fragment, no mate)."""
    rng = np.random.default_rng(seed)
    rng2 = np.random.default_rng(seed + 7919)      # draws of the optional extras: the plain shape of a seed never changes
    T = A.n_tx
    if multi_mean:
        n_fr = np.minimum(50, rng.geometric(1.0 / multi_mean, size=n_names)).astype(np.int64)
    else:
        n_fr = 1 + ((rng.random(n_names) < multi) * rng.integers(1, 4, size=n_names)).astype(np.int64)
    F = int(n_fr.sum())
    f_name = np.repeat(np.arange(n_names), n_fr)
    f_first = np.concatenate([[0], np.cumsum(n_fr)])[f_name]
    f_sec = (np.arange(F) - f_first) > 0
    tx0 = rng.integers(0, T, size=n_names)
    f_tx = tx0[f_name].copy()
    # secondary fragments: 70% another isoform of the same gene, else anywhere
    g_first = np.concatenate([[0], np.cumsum(np.bincount(A.t_gene))])     # transcripts of a gene are contiguous
    same = f_sec & (rng.random(F) < 0.7)
    g = A.t_gene[f_tx[same]]
    f_tx[same] = g_first[g] + (rng.random(int(same.sum())) * (g_first[g + 1] - g_first[g])).astype(np.int64)
    other = f_sec & ~same
    f_tx[other] = rng.integers(0, T, size=int(other.sum()))
    if getattr(A, "g_copies", None) is not None and int(A.g_copies.max()) > 1:
        par = np.nonzero(f_sec & (A.g_copies[A.t_gene[tx0[f_name]]] > 1))[0]      # the name's primary gene has paralogs
        t0 = tx0[f_name[par]]
        g0 = A.t_gene[t0]
        fam = A.g_family[g0]
        pick = A.fam_inst[A.fam_off[fam] + (rng2.random(par.size) * A.g_copies[g0]).astype(np.int64)]
        f_tx[par] = g_first[pick] + (t0 - g_first[g0])                           # same isoform, another locus
    L = A.t_len[f_tx]
    rl = np.minimum(read_len, L)
    frag = np.clip(rng.normal(300, 50, size=F).astype(np.int64), rl, L)
    a = (rng.random(F) * (L - frag + 1)).astype(np.int64)
    b = a + frag
    minus = A.t_minus[f_tx]
    if stranded == "fr":
        left_r1 = minus.copy()
    elif stranded == "rf":
        left_r1 = ~minus
    else:
        left_r1 = rng.random(F) < 0.5
    if stranded:
        flip = rng.random(F) < 0.03
        left_r1 = left_r1 ^ flip
    left_first = rng.random(F) < 0.5
    # alignments: 2 per fragment; slot 0/1 in file order
    N = 2 * F
    al_f = np.repeat(np.arange(F), 2)
    slot = np.tile(np.array([0, 1]), F)
    is_left = np.where(slot == 0, left_first[al_f], ~left_first[al_f])
    x = np.where(is_left, a[al_f], b[al_f] - rl[al_f])
    y = np.where(is_left, a[al_f] + rl[al_f], b[al_f])
    tx = f_tx[al_f]
    t_base = np.concatenate([[0], np.cumsum(A.t_len)])
    x_tx = np.repeat(np.arange(T), A.t_nex)
    gkey = t_base[x_tx] + A.x_cum
    k0 = np.searchsorted(gkey, t_base[tx] + x, side="right") - 1
    k1 = np.searchsorted(gkey, t_base[tx] + y - 1, side="right") - 1
    nseg = k1 - k0 + 1
    first_s = A.x_s[k0] + (x - A.x_cum[k0])
    last_e = A.x_s[k1] + (y - A.x_cum[k1])                     # exclusive
    lead = np.zeros(N, dtype=np.int64); trail = np.zeros(N, dtype=np.int64)
    joff = np.zeros(N, dtype=np.int64)
    interg = np.zeros(N, dtype=bool)
    if noise:
        r = rng.random(N); kk = rng.integers(1, 9, size=N); side = rng.random(N) < 0.5
        c1 = r < 0.02
        can_l = c1 & side & ((np.where(nseg == 1, last_e, A.x_e[k0] + 1) - first_s) > kk + 1)
        can_r = c1 & ~side & ((last_e - np.where(nseg == 1, first_s, A.x_s[k1])) > kk + 1)
        lead[can_l] = kk[can_l]; first_s[can_l] += kk[can_l]
        trail[can_r] = kk[can_r]; last_e[can_r] -= kk[can_r]
        c2 = (r >= 0.02) & (r < 0.035)
        first_s[c2 & side] -= kk[c2 & side]; last_e[c2 & ~side] += kk[c2 & ~side]
        c3 = (r >= 0.035) & (r < 0.045) & (nseg > 1)
        joff[c3] = rng.integers(1, 4, size=int(c3.sum())) * np.where(rng.random(int(c3.sum())) < 0.5, 1, -1)
        interg = (r >= 0.045) & (r < 0.065)
    # segments
    s_off = np.concatenate([[0], np.cumsum(nseg)])
    S = int(s_off[-1])
    s_al = np.repeat(np.arange(N), nseg)
    s_m = np.arange(S) - s_off[s_al]
    s_k = k0[s_al] + s_m
    seg_s = np.where(s_m == 0, first_s[s_al], A.x_s[s_k])
    seg_e = np.where(s_m == nseg[s_al] - 1, last_e[s_al], A.x_e[s_k] + 1)
    # junction offset on the first junction: move the end of segment 0 (stays > start)
    j0 = (s_m == 0) & (joff[s_al] != 0)
    ne = seg_e[j0] + joff[s_al][j0]
    ok = ne > seg_s[j0] + 1
    tmp = seg_e[j0]; tmp[ok] = ne[ok]; seg_e[j0] = tmp
    if interg.any():
        # collapse to one unspliced segment in the intergenic space before the transcript
        ia = np.nonzero(interg)[0]
        keep_seg = ~(interg[s_al] & (s_m > 0))
        seg_s[s_off[ia]] = np.maximum(1, A.t_start[tx[ia]] - 1400)
        seg_e[s_off[ia]] = seg_s[s_off[ia]] + rl[al_f[ia]]
        lead[ia] = 0; trail[ia] = 0
        seg_s, seg_e, s_al, s_m = seg_s[keep_seg], seg_e[keep_seg], s_al[keep_seg], s_m[keep_seg]
        nseg = nseg.copy(); nseg[ia] = 1
        s_off = np.concatenate([[0], np.cumsum(nseg)]); S = int(s_off[-1])
    # CIGAR words: [S] M (N M)* [S]
    ncig = 2 * nseg - 1 + (lead > 0) + (trail > 0)
    c_off = np.concatenate([[0], np.cumsum(ncig)])
    cig = np.zeros(int(c_off[-1]), dtype=np.uint32)
    has_lead = lead > 0
    cig[c_off[:-1][has_lead]] = (lead[has_lead] << 4) | 4
    has_tr = trail > 0
    cig[c_off[1:][has_tr] - 1] = (trail[has_tr] << 4) | 4
    mpos = c_off[:-1][s_al] + has_lead[s_al] + 2 * s_m
    cig[mpos] = ((seg_e - seg_s) << 4) | 0
    notfirst = s_m > 0
    gap = seg_s[notfirst] - seg_e[np.nonzero(notfirst)[0] - 1]
    cig[mpos[notfirst] - 1] = (gap << 4) | 3
    # flags / strand
    l_r1 = left_r1[al_f]
    sec = np.where(f_sec[al_f], 256, 0)
    flag = np.where(is_left, 1 | 2 | 32 | np.where(l_r1, 64, 128), 1 | 2 | 16 | np.where(l_r1, 128, 64)) | sec
    spliced = nseg > 1
    xs = np.zeros(N, dtype=np.int8)
    if stranded:
        strand = np.where(l_r1, ord("-"), ord("+")) if stranded == "fr" else np.where(l_r1, ord("+"), ord("-"))
    else:
        has_xs = spliced & (rng.random(N) < 0.9)
        xs[has_xs] = np.where(minus[al_f][has_xs], ord("-"), ord("+"))
        strand = np.where(has_xs, xs, ord("."))
    B = FastBatch()
    pos0 = seg_s[s_off[:-1]] - 1
    mate = np.arange(N) ^ 1
    grp = np.concatenate([[0], np.cumsum(2 * n_fr)])
    refid = ix.refid_of_chrom[A.t_chrom[tx]]
    al_name, chrom, rlen = f_name[al_f] + name_base, A.t_chrom[tx], rl[al_f].astype(np.int64)
    if paired_frac < 1.0:
        # single-end names keep the first alignment of every fragment: no PAIRED/PROPER/MREVERSE/READ1/READ2 bits, no mate
        single = rng2.random(n_names) >= paired_frac
        keep = ~(single[f_name[al_f]] & (slot == 1))
        flag = np.where(single[f_name[al_f]], flag & (16 | 256), flag)
        new_idx = np.cumsum(keep) - 1
        mate = np.where(single[f_name[al_f]], -1, new_idx[mate])[keep]
        kseg, kcig = keep[s_al], np.repeat(keep, ncig)
        seg_s, seg_e, cig = seg_s[kseg], seg_e[kseg], cig[kcig]
        s_off = np.concatenate([[0], np.cumsum(nseg[keep])]); c_off = np.concatenate([[0], np.cumsum(ncig[keep])])
        grp = np.concatenate([[0], np.cumsum(np.where(single, n_fr, 2 * n_fr))])
        refid, strand, flag, pos0, xs = refid[keep], strand[keep], flag[keep], pos0[keep], xs[keep]
        al_name, chrom, rlen = al_name[keep], chrom[keep], rlen[keep]
        N = int(keep.sum())
    B.finish(refid, strand, flag, s_off, seg_s, seg_e, c_off, cig, mate, grp, None, None, np.arange(N), None)
    B.flag = flag.astype(np.int64); B.pos0 = pos0; B.al_name = al_name; B.chrom = chrom
    B.read_len = rlen; B.xs = xs; B.n_names = n_names
    return B


def to_records(B, lo_name=None, hi_name=None, seed=0):
    """BAM record dicts (synth.write_bam format) for the alignments of groups [lo_name, hi_name)."""
    rng = np.random.default_rng(seed)
    A = B.arrays
    g0 = 0 if lo_name is None else lo_name
    g1 = B.desc.n_groups if hi_name is None else hi_name
    a0, a1 = int(A["group_off"][g0]), int(A["group_off"][g1])
    ops = "MIDNSHP=X"
    recs = []
    co = A["cigar_off"].astype(np.int64)
    cg = A["cigar"][int(co[a0]):int(co[a1])]
    cw_op, cw_len = (cg & 15).tolist(), (cg >> 4).tolist()
    qcons = np.isin(cg & 15, (0, 1, 4, 7, 8)) * (cg >> 4).astype(np.int64)          # M I S = X consume the read
    qsum = np.concatenate([[0], np.cumsum(qcons)])
    ql_of = (qsum[co[a0 + 1:a1 + 1] - co[a0]] - qsum[co[a0:a1] - co[a0]]).tolist()
    bases = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=int(qsum[-1]))].tobytes()
    mate = A["mate_idx"].tolist()
    flag, chrom, pos0, xs, name = B.flag.tolist(), B.chrom.tolist(), B.pos0.tolist(), B.xs.tolist(), B.al_name.tolist()
    quals = {}
    bp = 0
    for a in range(a0, a1):
        m = mate[a] if mate[a] >= 0 else a
        c0, c1 = int(co[a] - co[a0]), int(co[a + 1] - co[a0])
        cigar = [(ops[cw_op[i]], cw_len[i]) for i in range(c0, c1)]
        ql = ql_of[a - a0]
        tags = [("NH", "C", 1), ("HI", "C", 1), ("AS", "C", 200), ("NM", "C", 0)]
        if xs[a]:
            tags.insert(2, ("XS", "A", chr(xs[a])))
        if ql not in quals:
            quals[ql] = bytes([30]) * ql
        recs.append(dict(name=f"r{name[a]:09d}", flag=flag[a], chrom=chrom[a], pos=pos0[a], mapq=255,
                         cigar=cigar, mchrom=chrom[m] if mate[a] >= 0 else -1, mpos=pos0[m] if mate[a] >= 0 else -1, isize=0,
                         seq=bases[bp:bp + ql], qual=quals[ql], tags=tags))
        bp += ql
    return recs
