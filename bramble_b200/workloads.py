"""The five synthetic configurations of BASELINE.json / SURVEY.md 8(d) at bench scale, as engine inputs.

    c1  1 chromosome, 500 genes x ~4 isoforms (~2k transcripts); 1M 150 bp pairs, default mode (unstranded search)
    c2  GENCODE-shaped annotation (~250k transcripts, 25 chromosomes); --fr pairs      <- the configuration the metric is quoted on
    c3  C1-shaped x 10 chromosomes + random genome; ONT-like long reads (8 % errors, soft clips), --lr -S (ksw2 clip rescue)
    c4  human-scale annotation + 200 genes with 50-200 isoforms and 150-300 exons; HiFi-like long reads, --lr-hq --strict
    c5  C2 annotation + 2000 paralog families at 2-50 loci; 1+Geometric(5) alignments per name, 60 % paired, default mode

`build(name, seed, names)` returns everything a consumer needs: index, bundle (structure-of-arrays), resolved thresholds,
the reference's command-line flags, and writers for the files the reference arm reads (GTF, FASTA, a BAM of the first
names of the same bundle).  The full configs are 10^7-10^8 reads; `names` per GPU is sized so that one bundle is at
least the reference's bundle size (>= 100 000 alignments, src/bramble.cpp:362) and the default bench run ends in minutes.
"""
import os

import numpy as np

from . import capi, synth, synth_fast as sf, synth_par

DEFAULT_NAMES = {"c1": 1_000_000, "c2": 2_000_000, "c3": 100_000, "c4": 60_000, "c5": 400_000}
FLAGS = {"c1": [], "c2": ["--fr"], "c3": ["--lr"], "c4": ["--lr-hq", "--strict"], "c5": []}      # c3 also gets -S <fasta>
LONG = {"c3", "c4"}
SHAPE = {
    "c1": "C1: 1 chromosome, 500 genes / ~2k transcripts + unstranded paired-end 150bp name-collated reads, default mode",
    "c2": "C2: synthetic GENCODE-shaped GTF (252,687 transcripts, 2.24M exons, 25 chromosomes) + --fr paired-end 150bp name-collated reads",
    "c3": "C3: C1-shaped annotation x 10 chromosomes (~19k transcripts) + 190 Mb random genome + ONT-like long reads (median 1.2 kb, 8% I/D/X errors, "
          "50% soft-clipped ends, 70% of the clips true neighbour-exon sequence), --lr -S: INS/GAP segments, similarity filter, ksw2 clip rescue",
    "c4": "C4: GENCODE-shaped GTF + 200 genes with 50-200 isoforms and 150-300 exons (uint8 exon_id wrap) + HiFi-like long reads (median 2.5 kb, "
          "0.5% errors), --lr-hq --strict",
    "c5": "C5: GENCODE-shaped GTF with 2000 paralog families at 2-50 loci + unstranded 150bp reads, 1+Geometric(5) alignments per read name "
          "(cap 50, secondaries on the family's other loci), 60% of the names paired, default mode",
}


class Workload:
    pass


def _annotation(name):
    if name == "c1":
        return sf.make_annotation(1001, n_chrom=1, n_genes=500, iso_mean=4.0, iso_cap=12)
    if name == "c2":
        return sf.make_annotation(2002)
    if name == "c3":
        return sf.make_annotation(3003, n_chrom=10, n_genes=5000, iso_mean=4.0, iso_cap=12)
    if name == "c4":
        return sf.make_annotation(4004, hi_iso_genes=200, hi_iso_exons=300)
    if name == "c5":
        return sf.make_annotation(5005, paralog_families=2000, paralog_cap=50)
    raise ValueError(name)


def build(name, seed=1000, names=None, sample_names=0):
    """sample_names: long-read workloads keep the BAM records of (at least) that many leading names for `write_sample`."""
    names = names or DEFAULT_NAMES[name]
    W = Workload()
    W.name, W.names, W.flags, W.shape, W.long = name, names, list(FLAGS[name]), SHAPE[name], name in LONG
    W.A = A = _annotation(name)
    W.genome = sf.make_genome_codes(A, 3004) if name == "c3" else None
    W.ix = sf.build_index(A, genome_codes=W.genome)
    W.recs = None
    if name == "c1":
        W.B = sf.make_short_batch(A, W.ix, seed, names, stranded=None, multi=0.0)
        W.params = capi.resolve_params()
    elif name == "c2":
        W.B = sf.make_short_batch(A, W.ix, seed, names, stranded="fr")
        W.params = capi.resolve_params()
    elif name == "c5":
        W.B = sf.make_short_batch(A, W.ix, seed, names, stranded=None, multi_mean=5.0, paired_frac=0.6)
        W.params = capi.resolve_params()
    else:
        ann = sf.to_annotation(A)
        chunk = 2000 if name == "c3" else 3000
        keep = (sample_names + chunk - 1) // chunk
        if name == "c3":
            W.B, W.recs = synth_par.make_long_batch(ann, sf.genome_bytes(W.genome), W.ix, seed, names, params=dict(err=0.08), median_len=1200,
                                                    chunk=chunk, keep_records=keep)
            W.params = capi.resolve_params(lr=True, use_fasta=True)
        else:
            W.B, W.recs = synth_par.make_long_batch(ann, None, W.ix, seed, names, params=dict(err=0.005), median_len=2500,
                                                    chunk=chunk, keep_records=keep)
            W.params = capi.resolve_params(lr_hq=True, strict=True)
    return W


def write_annotation(W, d):
    """GTF (+ FASTA for c3) the reference arm reads; returns the extra command-line arguments."""
    sf.write_gtf(W.A, os.path.join(d, "a.gtf"))
    if W.genome is not None:
        sf.write_fasta(W.A, W.genome, os.path.join(d, "g.fa"))
        return ["-S", os.path.join(d, "g.fa")]
    return []


def write_decoy_annotation(W, d):
    """The same GTF on renamed chromosomes: no read finds a transcript, so a run against it pays BAM decode, ingest and
    bundling but no projection (the `decode-only pass` of SURVEY 8(d))."""
    with open(os.path.join(d, "a.gtf")) as f, open(os.path.join(d, "decoy.gtf"), "w") as g:
        for line in f:
            g.write("zz_" + line)


def write_sample(W, path, n_names):
    """BAM of the first n_names read names of the bundle; returns the number of (mapped) alignments written."""
    class Ann:
        chroms, chrom_len = W.A.chroms, W.A.chrom_len
    if W.long:
        recs, seen, prev = [], 0, None
        for r in W.recs:
            if r["name"] != prev:
                if seen == n_names:
                    break
                seen += 1
                prev = r["name"]
            recs.append(r)
    else:
        recs = sf.to_records(W.B, 0, min(n_names, int(W.B.desc.n_groups)))
    synth.write_bam(path, Ann, recs)
    return sum(1 for r in recs if not r["flag"] & 4)
