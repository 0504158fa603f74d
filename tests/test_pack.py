"""CPU: host-side ingest logic (Python mirror of the reference's pre-kernel half)."""
import numpy as np

from bramble_b200 import pack, shard
import common


def w(s):
    import re
    return [(int(n) << 4) | "MIDNSHP=X".index(o) for n, o in re.findall(r"(\d+)([MIDNSHP=X])", s)]


def test_read_exons_follow_gsam_setup_coordinates():
    assert pack.read_exons(100, w("50M")) == [(101, 151)]
    assert pack.read_exons(100, w("5S20M100N30M3S")) == [(101, 121), (221, 251)]
    assert pack.read_exons(100, w("10M2I10M3D10M")) == [(101, 134)]          # I ignored, D extends
    assert pack.read_exons(100, w("100N20M")) == [(101, 121)]                 # leading N before any M is ignored
    assert pack.read_exons(100, w("10M50N5I50N10M")) == [(101, 111), (211, 221)]   # I between two N's suppresses the close
    assert pack.read_exons(100, w("10M50N")) == [(101, 111)]                  # trailing intron: no final exon
    assert pack.read_exons(100, w("10=1X10=")) == [(101, 122)]


def test_strand_rules():
    r = dict(flag=0, tags=[("XS", "A", "-")])
    assert pack.read_strand(r) == "-"
    assert pack.read_strand(dict(flag=16, tags=[("ts", "A", "+")])) == "-"    # minimap2 ts flipped by FREVERSE
    assert pack.read_strand(dict(flag=0, tags=[("ts", "A", "+")])) == "+"
    assert pack.read_strand(dict(flag=0, tags=[])) == "."
    # --fr: read1 forward -> '-', read2 forward -> '+'; unpaired forward -> '-'
    assert pack.read_strand(dict(flag=1 | 0x40, tags=[]), fr=True) == "-"
    assert pack.read_strand(dict(flag=1 | 0x40 | 16, tags=[]), fr=True) == "+"
    assert pack.read_strand(dict(flag=1 | 0x80, tags=[]), fr=True) == "+"
    assert pack.read_strand(dict(flag=0, tags=[]), fr=True) == "-"
    assert pack.read_strand(dict(flag=1 | 0x40 | 16, tags=[]), rf=True) == "-"


def test_mates_groups_and_bundles():
    wl = common.make_workload("default", 5, n_names=600, n_genes=20, chunk_size=150)
    assert len(wl["batches"]) > 2
    total = 0
    for b in wl["batches"]:
        A = b.arrays
        n = b.desc.n_aln
        total += n
        m = A["mate_idx"]
        for a in range(n):
            if m[a] >= 0:
                assert m[m[a]] == a and m[a] != a                      # single, symmetric links
                assert b.names[a] == b.names[m[a]]
        go = A["group_off"]
        assert go[0] == 0 and go[-1] == n
        for g in range(len(go) - 1):
            names = {b.names[a] for a in range(go[g], go[g + 1])}
            assert len(names) == 1
        if b is not wl["batches"][-1]:
            assert n >= 150                                             # cut only at a name change after >= chunk alignments
    assert total == sum(1 for r in wl["recs"] if not r["flag"] & 4)


def test_sort_perm_is_a_permutation_sorted_by_start():
    rng = np.random.default_rng(1)
    st = rng.integers(0, 50, size=5000).astype(np.int32)
    perm = pack.sort_perm_by_start(st)
    assert sorted(perm.tolist()) == list(range(5000))
    assert np.all(np.diff(st[perm]) >= 0)


def test_shard_dealing():
    assert shard.bundles_of_rank(7, 2, 0) == [0, 2, 4, 6] and shard.bundles_of_rank(7, 2, 1) == [1, 3, 5]
    assert [shard.owner_of(k, 4) for k in range(6)] == [0, 1, 2, 3, 0, 1]


def test_hostbind_cpulist_and_no_gpu():
    """Host placement helper: cpulist parsing, and a box without NUMA information (or without a GPU) is left untouched."""
    import os
    from bramble_b200 import hostbind
    assert hostbind._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert hostbind._parse_cpulist("") == set()
    before = os.sched_getaffinity(0)
    info = hostbind.bind_to_gpu_node(0)
    if info["numa_node"] is None:
        assert not info["bound"] and os.sched_getaffinity(0) == before
    os.sched_setaffinity(0, before)


def test_hostbind_even_slices_without_numa_information():
    """One process per GPU on a box that exposes no NUMA node for the device: rank r of n takes the r-th even slice of
    the allowed CPUs, slices are disjoint, and a box with fewer than two CPUs per rank is left alone."""
    import os
    from bramble_b200 import hostbind
    before = os.sched_getaffinity(0)
    try:
        if hostbind.gpu_numa_node(0) is not None or len(before) < 4:
            return
        seen = []
        for r in range(2):
            os.sched_setaffinity(0, before)
            info = hostbind.bind_to_gpu_node(0, r, 2)
            got = os.sched_getaffinity(0)
            assert info["bound"] and info["cpus_allowed"] == len(got) == len(before) // 2
            seen.append(got)
        assert not (seen[0] & seen[1]) and (seen[0] | seen[1]) <= before
        os.sched_setaffinity(0, before)
        info = hostbind.bind_to_gpu_node(0, 0, len(before))          # one CPU per rank: not worth pinning
        assert not info["bound"] and os.sched_getaffinity(0) == before
    finally:
        os.sched_setaffinity(0, before)


def test_compact_columns_say_the_same_bundle():
    """Batch.compact(): the 16-bit counts / links and the position are the plain columns in fewer bytes -- the offsets are the running sums
    of the counts, the links resolve to the same mates, and the read exons follow from position + CIGAR by the ingest rule (pack.read_exons,
    which the engine restates on the device) -- for reference-shaped bundles (pack_batches) and for the bench generator's (synth_fast)."""
    from bramble_b200 import pack, synth_fast as sf
    import common
    batches = []
    for mode in ("default", "lr"):
        batches += common.make_workload(mode, 151, n_names=600, n_genes=30)["batches"]
    A = sf.make_annotation(77, n_chrom=2, n_genes=120)
    batches.append(sf.make_short_batch(A, sf.build_index(A), 78, 3000, stranded="fr"))
    for b in batches:
        n = int(b.desc.n_aln)
        c = b.compact()
        X, Y = b.arrays, c.arrays
        assert c.desc.n_aln == n and int(c.desc.n_cigar_total) == int(X["cigar_off"][n])
        assert not c.desc.exon_off and not c.desc.cigar_off and not c.desc.mate_idx and not c.desc.refid          # the replaced columns are NULL
        assert np.array_equal(np.concatenate([[0], np.cumsum(Y["cigar_cnt16"][:n].astype(np.int64))]), X["cigar_off"][:n + 1].astype(np.int64))
        mate = np.where(Y["mate_delta16"][:n] != 0, np.arange(n) + Y["mate_delta16"][:n].astype(np.int64), -1)
        assert np.array_equal(mate, X["mate_idx"][:n].astype(np.int64)) and np.array_equal(Y["refid16"][:n].astype(np.int64), X["refid"][:n].astype(np.int64))
        co, eo = X["cigar_off"].astype(np.int64), X["exon_off"].astype(np.int64)
        for a in range(0, n, max(1, n // 400)):
            ex = pack.read_exons(int(Y["pos0"][a]), [int(w) for w in X["cigar"][co[a]:co[a + 1]]])
            assert [s for s, _ in ex] == list(X["exon_st"][eo[a]:eo[a + 1]]) and [e for _, e in ex] == list(X["exon_en"][eo[a]:eo[a + 1]]), a
        # (a long-read bundle is CIGAR words: little to gain there; short reads lose a third to a half of their bytes)
        assert sum(v.nbytes for k, v in Y.items() if hasattr(c.desc, k)) < sum(v.nbytes for k, v in X.items() if hasattr(b.desc, k))
