"""Shared builders for the tests: seeded synthetic workloads per mode."""
import numpy as np

from bramble_b200 import capi, pack, synth

MODES = {
    "default": dict(args=[], long=False),
    "fr": dict(args=["--fr"], long=False, fr=True),
    "rf": dict(args=["--rf"], long=False, rf=True),
    "strict": dict(args=["--strict"], long=False, strict=True),
    "lr": dict(args=["--lr"], long=True, lr=True, genome=True),
    "lr_nofasta": dict(args=["--lr"], long=True, lr=True),
    "lrhq": dict(args=["--lr-hq", "--strict"], long=True, lr_hq=True, strict=True, hi_iso=True),
}


def params_for(mode):
    m = MODES[mode]
    return capi.resolve_params(lr=m.get("lr", False), lr_hq=m.get("lr_hq", False), strict=m.get("strict", False),
                               use_fasta=m.get("genome", False))


def make_workload(mode, seed, n_names=2000, n_genes=50, n_chrom=2, chunk_size=100000, micro_introns=False):
    m = MODES[mode]
    ann = synth.make_annotation(seed, n_chrom=n_chrom, n_genes=n_genes, hi_iso_genes=2 if m.get("hi_iso") else 0, hi_iso_exons=300)
    if micro_introns:
        add_micro_introns(ann, seed)
    genome = synth.make_genome(ann, seed + 1) if m.get("genome") else None
    if m["long"]:
        recs = synth.make_long_reads(ann, genome, seed + 2, n_names, params=dict(err=0.005 if m.get("lr_hq") else 0.08))
    else:
        recs = synth.make_short_reads(ann, seed + 2, n_names, stranded="fr" if m.get("fr") else ("rf" if m.get("rf") else None))
    ix = pack.build_index(ann, genome=genome)
    batches = pack.pack_batches(recs, ix, ann.chroms, long_reads=m["long"], fr=m.get("fr", False), rf=m.get("rf", False),
                                use_fasta=genome is not None, chunk_size=chunk_size)
    return dict(ann=ann, genome=genome, recs=recs, index=ix, batches=batches, params=params_for(mode), mode=m)


def add_micro_introns(ann, seed):
    """Split some exons by a 1-30 bp intron (real annotations have them): one read exon can then hit two
    guide exons of the same transcript (duplicate-tid paths of src/evaluate.cpp:218-224 and the
    injectivity check :1031-1052)."""
    rng = np.random.default_rng(seed + 99)
    for t in ann.tx:
        if rng.random() < 0.3:
            ex = list(t["exons"])
            k = int(rng.integers(len(ex)))
            s, e = ex[k]
            if e - s > 80:
                cut = int(rng.integers(s + 30, e - 40))
                gap = int(rng.integers(1, 31))
                ex[k:k + 1] = [(s, cut), (cut + gap + 1, e)]
                t["exons"] = ex
