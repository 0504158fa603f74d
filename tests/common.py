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


def ksw_problems(seed=17, n=400, short_max=120):
    """Random extension problems for ksw2 (codes 0..4): every 4th an unrelated target, the others the query with
    substitutions / indels / Ns plus a tail (as a clip against its neighbour exon looks); every 10th a long one
    (300-900 bases, beyond the kernels' shared-memory state); z-drop and `score == NEG_INF` both occur."""
    rng = np.random.default_rng(seed)
    qs, ts = [], []
    for k in range(n):
        ql = int(rng.integers(1, short_max)) if k % 10 else int(rng.integers(300, 900))
        q = rng.integers(0, 4, size=ql).astype(np.uint8)
        kind = k % 4
        if kind == 0:       # unrelated target
            t = rng.integers(0, 4, size=int(rng.integers(1, ql + 41))).astype(np.uint8)
        else:               # target = query with substitutions / indels / Ns, plus a tail
            t = list(q)
            for _ in range(int(rng.integers(0, max(2, ql // 12)))):
                pos = int(rng.integers(0, len(t)))
                r = rng.random()
                if r < 0.4: t[pos] = int(rng.integers(0, 5))
                elif r < 0.7: del t[pos]
                else: t.insert(pos, int(rng.integers(0, 4)))
                if not t: t = [0]
            t = np.array(t + list(rng.integers(0, 4, size=int(rng.integers(0, 41)))), dtype=np.uint8)[:ql + 40]
        qs.append(q); ts.append(t)
    return qs, ts
