#!/usr/bin/env python
"""Generates tests/golden/annot/ -- realistic annotation files for the loader differential:

  gencode_like.gtf   GENCODE-style GTF: gene / transcript / exon / CDS / UTR / start_codon / stop_codon rows, versioned IDs,
                     attributes in varying order, ';' inside a quoted value, '-' strand exons listed in transcript (descending)
                     order, chromosomes in non-alphabetical file order, a transcript without a transcript row, a transcript
                     row without exon rows, overlapping and adjacent exon rows (gclib merges them), interleaved transcripts
  ensembl_like.gff3  the same models as GFF3: ID= / Parent= (an exon shared by two parents), gene and CDS rows, mRNA and
                     transcript features
  reads.bam          short reads over those models (name-collated pairs and singles)
  expected_*.bam     what the UNMODIFIED reference (oracle/_ref/bramble_ref) writes for each annotation

The tests (tests/test_annotations.py) run the drop-in CLI on the committed files and compare the decompressed streams:
@SQ order and lengths, record order, every record byte.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from bramble_b200 import synth  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "bramble_ref")
OUT = os.path.join(HERE, "annot")

# (transcript id, gene id, chrom, strand, exons as written (may overlap / touch / be unordered), cds span or None, style)
MODELS = [
    ("ENST00000000101.3", "ENSG00000000011.7", "chr2", "+", [(1000, 1200), (2000, 2150), (3000, 3400)], (1100, 3100), "full"),
    ("ENST00000000102.1", "ENSG00000000011.7", "chr2", "+", [(1000, 1200), (3000, 3400)], (1100, 3100), "full"),
    ("ENST00000000103.2", "ENSG00000000011.7", "chr2", "+", [(950, 1200), (2000, 2150), (2151, 2300), (3000, 3300)], None, "full"),      # adjacent exon rows -> merged
    ("ENST00000000201.5", "ENSG00000000022.2", "chr2", "-", [(8000, 8200), (9000, 9100), (9050, 9300)], (8100, 9200), "full"),          # overlapping exon rows -> merged
    ("ENST00000000202.1", "ENSG00000000022.2", "chr2", "-", [(8000, 8200), (8600, 8650), (9000, 9300)], None, "noline"),               # no transcript row
    ("ENST00000000301.1", "ENSG00000000033.1", "chr10", "+", [(500, 900)], None, "noexon"),                                             # transcript row only: implicit exon
    ("ENST00000000302.4", "ENSG00000000033.1", "chr10", "+", [(500, 700), (1200, 1500), (1900, 2400)], (600, 2000), "full"),
    ("ENST00000000401.2", "ENSG00000000044.9", "chr1", "-", [(300, 450), (700, 800), (1000, 1400)], (350, 1300), "full"),
    ("ENST00000000402.2", "ENSG00000000044.9", "chr1", "-", [(300, 450), (1000, 1400)], None, "full"),
    ("ENST00000000403.1", "ENSG00000000044.9", "chr1", "+", [(300, 450), (1000, 1400)], None, "full"),                                   # same coordinates, other strand
    ("ENST00000000501.1", "ENSG00000000055.3", "chrX_alt", "+", [(10, 120), (200, 260)], None, "full"),
]
CHROM_LEN = {"chr2": 20000, "chr10": 5000, "chr1": 5000, "chrX_alt": 1000}
FILE_CHROM_ORDER = ["chr2", "chr10", "chr1", "chrX_alt"]           # refid = first appearance; tids follow the NAME order chr1 < chr10 < chr2 < chrX_alt


def merged(exons):
    out = []
    for s, e in sorted(exons):
        if out and s <= out[-1][1] + 1:
            out[-1][1] = max(out[-1][1], e)
        else:
            out.append([s, e])
    return [tuple(x) for x in out]


def write_gtf(path):
    L = ["##description: synthetic GENCODE-shaped annotation for the bramble_b200 loader differential", "##provider: tests/golden/make_annotations.py",
         "##format: gtf"]
    genes = {}
    for m in MODELS:
        genes.setdefault(m[1], []).append(m)
    done_gene = set()
    pending = []                                   # exon rows of the previous transcript that are written after the next transcript's row (interleaving)
    for k, (tid, gid, chrom, strand, exons, cds, style) in enumerate(MODELS):
        gname = "GENE; %d" % k                     # a ';' inside a quoted value
        if gid not in done_gene:
            done_gene.add(gid)
            ms = genes[gid]
            gs, ge = min(e[0] for m in ms for e in m[4]), max(e[1] for m in ms for e in m[4])
            L.append(f'{chrom}\tHAVANA\tgene\t{gs}\t{ge}\t.\t{ms[0][3]}\t.\tgene_id "{gid}"; gene_type "protein_coding"; gene_name "{gname}"; level 2;')
        ts, te = min(e[0] for e in exons), max(e[1] for e in exons)
        head = f'gene_id "{gid}"; transcript_id "{tid}"; gene_type "protein_coding"; gene_name "{gname}"; transcript_type "protein_coding"; transcript_name "T-{k}";'
        alt = f'transcript_id "{tid}"; level 2; gene_id "{gid}"; tag "basic"; gene_name "{gname}";'      # other attribute order
        if style != "noline":
            L.append(f'{chrom}\tHAVANA\ttranscript\t{ts}\t{te}\t.\t{strand}\t.\t{head} level 2; tag "basic"; tag "Ensembl_canonical";')
        L.extend(pending)
        pending = []
        if style == "noexon":
            continue
        rows = sorted(exons, reverse=(strand == "-"))                 # GENCODE lists '-' strand exons in transcript order
        for n, (s, e) in enumerate(rows):
            row = f'{chrom}\tHAVANA\texon\t{s}\t{e}\t.\t{strand}\t.\t{alt if n % 2 else head} exon_number {n + 1}; exon_id "ENSE{k:05d}{n:03d}.1";'
            if k % 3 == 0 and n == len(rows) - 1:
                pending.append(row)                                   # the last exon row of this transcript comes after the next transcript row
            else:
                L.append(row)
            if cds and s <= cds[1] and e >= cds[0]:
                cs, ce = max(s, cds[0]), min(e, cds[1])
                L.append(f'{chrom}\tHAVANA\tCDS\t{cs}\t{ce}\t.\t{strand}\t0\t{head} exon_number {n + 1}; protein_id "ENSP{k:05d}.1";')
                if cs > s:
                    L.append(f'{chrom}\tHAVANA\tUTR\t{s}\t{cs - 1}\t.\t{strand}\t.\t{head} exon_number {n + 1};')
        if cds:
            L.append(f'{chrom}\tHAVANA\tstart_codon\t{cds[0]}\t{cds[0] + 2}\t.\t{strand}\t0\t{head}')
            L.append(f'{chrom}\tHAVANA\tstop_codon\t{cds[1] - 2}\t{cds[1]}\t.\t{strand}\t0\t{head}')
    L.extend(pending)
    open(path, "w").write("\n".join(L) + "\n")


def write_gff3(path):
    L = ["##gff-version 3"] + [f"##sequence-region {c} 1 {CHROM_LEN[c]}" for c in FILE_CHROM_ORDER]
    done_gene = set()
    genes = {}
    for m in MODELS:
        genes.setdefault(m[1], []).append(m)
    shared = {}                                                      # an exon row listing two parents: same gene, same coordinates, written once
    for k, (tid, gid, chrom, strand, exons, cds, style) in enumerate(MODELS):
        if gid not in done_gene:
            done_gene.add(gid)
            ms = genes[gid]
            gs, ge = min(e[0] for m in ms for e in m[4]), max(e[1] for m in ms for e in m[4])
            L.append(f"{chrom}\tensembl\tgene\t{gs}\t{ge}\t.\t{ms[0][3]}\t.\tID=gene:{gid};Name=G{k};biotype=protein_coding;description=a gene%3B with an escaped semicolon")
        ts, te = min(e[0] for e in exons), max(e[1] for e in exons)
        feat = "mRNA" if k % 2 == 0 else "transcript"
        L.append(f"{chrom}\tensembl\t{feat}\t{ts}\t{te}\t.\t{strand}\t.\tID=transcript:{tid};Parent=gene:{gid};Name=T{k};biotype=protein_coding;tag=basic")
        if style == "noexon":
            continue
        for n, (s, e) in enumerate(sorted(exons)):
            key = (gid, chrom, strand, s, e)
            if k in (0, 1) and (s, e) == (1000, 1200):               # models 0 and 1 share their first exon: one row, two parents
                if key in shared:
                    continue
                shared[key] = True
                L.append(f"{chrom}\tensembl\texon\t{s}\t{e}\t.\t{strand}\t.\tParent=transcript:{MODELS[0][0]},transcript:{MODELS[1][0]};Name=ENSE{k:05d}{n:03d};rank={n + 1}")
                continue
            L.append(f"{chrom}\tensembl\texon\t{s}\t{e}\t.\t{strand}\t.\tParent=transcript:{tid};Name=ENSE{k:05d}{n:03d};constitutive=0;rank={n + 1}")
            if cds and s <= cds[1] and e >= cds[0]:
                L.append(f"{chrom}\tensembl\tCDS\t{max(s, cds[0])}\t{min(e, cds[1])}\t.\t{strand}\t0\tID=CDS:ENSP{k:05d};Parent=transcript:{tid}")
    open(path, "w").write("\n".join(L) + "\n")


def make_reads():
    """Reads over the MERGED models (what the annotation means), via the synthetic generator."""
    chroms = FILE_CHROM_ORDER
    tx = [dict(id=tid, gene=gid, chrom=chroms.index(chrom), strand=strand, exons=merged(exons), family=i) for i, (tid, gid, chrom, strand, exons, cds, style) in enumerate(MODELS)]
    ann = synth.Annotation(chroms, [CHROM_LEN[c] for c in chroms], tx)
    recs = synth.make_short_reads(ann, 2024, 900, read_len=100)
    return ann, recs


def main():
    assert os.path.exists(REF), "build the reference first: make -C oracle ref"
    shutil.rmtree(OUT, ignore_errors=True)
    os.makedirs(OUT)
    write_gtf(os.path.join(OUT, "gencode_like.gtf"))
    write_gff3(os.path.join(OUT, "ensembl_like.gff3"))
    ann, recs = make_reads()
    synth.write_bam(os.path.join(OUT, "reads.bam"), ann, recs)
    for name in ("gencode_like.gtf", "ensembl_like.gff3"):
        wd = os.path.join(OUT, "run")
        os.makedirs(wd, exist_ok=True)
        argv = ["bramble", "../reads.bam", "-G", "../" + name, "-o", "out.bam", "-p", "1", "--quiet"]
        r = subprocess.run(argv, executable=REF, cwd=wd, capture_output=True, text=True, env=dict(os.environ, BRB_SHIM_LEVEL="6"))
        assert r.returncode == 0, r.stderr
        os.replace(os.path.join(wd, "out.bam"), os.path.join(OUT, "expected_" + name.split(".")[0] + ".bam"))
        shutil.rmtree(wd)
        text, refs, out = synth.read_bam(os.path.join(OUT, "expected_" + name.split(".")[0] + ".bam"))
        print(name, "->", len(out), "records; @SQ:", [r_[0] + ":" + str(r_[1]) for r_ in refs])


if __name__ == "__main__":
    main()
