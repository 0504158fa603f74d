"""GPU parity: the CUDA engine (through the C ABI) against the CPU oracle on the same seeded inputs.
Bit-exact for every integer column and CIGAR; similarity scores bit-exact (tolerance 1e-6 relative is
the north_star fallback, asserted in oracle_lib.assert_rows_equal)."""
import numpy as np
import pytest

from bramble_b200 import capi, engine
import common
import oracle_lib

pytestmark = pytest.mark.gpu


def _check(wl, what):
    ix, p = wl["index"], wl["params"]
    n_rec = 0
    with engine.Engine(ix) as eng:
        for bi, b in enumerate(wl["batches"]):
            exp = oracle_lib.port_project(ix, p, b)
            alg = None
            # every front kernel: the default choice (short-read fast-path kernel where it applies), the short-read
            # kernel without the fast path, and the generic one; rows as unit blocks (default), in emit order, and as
            # 16-byte packed records, 12-byte or 8-byte compact ones (long-read modes fall back to the 16-byte records)
            for generic, nofast, ordered, packed in ((0, 0, 0, 0), (0, 1, 0, 0), (1, 0, 0, 0), (0, 0, 1, 0), (0, 0, 0, 1), (1, 0, 0, 1), (0, 0, 0, 2), (1, 0, 0, 2), (0, 0, 0, 3), (0, 1, 0, 3)):
                eng.set_option(capi.OPT_GENERIC_FRONT, generic)
                eng.set_option(capi.OPT_NO_FASTPATH, nofast)
                eng.set_option(capi.OPT_ORDERED_ROWS, ordered)
                eng.set_option(capi.OPT_PACKED_ROWS, packed)
                got = eng.project(p, b)
                oracle_lib.assert_rows_equal(exp, got, f"{what} batch {bi} generic={generic} nofast={nofast} ordered={ordered} packed={packed}")
                st = eng.stats()
                assert st.n_rec == got["n_rec"] and st.n_launches > 0
                if generic == 0 and nofast == 0:          # the work counters do not depend on the row layout
                    alg = alg if (ordered or packed) else int(st.alg_bytes)
                    assert int(st.alg_bytes) == alg, (ordered, packed)
                assert st.n_matches == exp["n_matches"]
                assert st.n_rescue == exp["n_rescue"]
            n_rec += got["n_rec"]
    assert n_rec > 0
    return n_rec


@pytest.mark.parametrize("mode", ["default", "fr", "rf", "strict", "lr", "lr_nofasta", "lrhq"])
@pytest.mark.parametrize("seed", [21, 22])
def test_project_matches_oracle(mode, seed):
    wl = common.make_workload(mode, seed, n_names=1500 if common.MODES[mode]["long"] else 3000)
    _check(wl, f"{mode}/{seed}")


@pytest.mark.parametrize("mode", ["default", "lr", "lr_nofasta", "lrhq"])
def test_micro_introns_duplicate_tids(mode):
    wl = common.make_workload(mode, 31, n_names=2000, micro_introns=True)
    _check(wl, f"micro/{mode}")


@pytest.mark.parametrize("mode", ["default", "lr_nofasta"])
def test_hot_genes_big_batches(mode):
    """Genes with 50-200 isoforms: a batch of 32 alignments then carries far more matches than the back kernel
    stages in shared memory (global staging, general CIGAR merges on the second walk) and the front kernels walk
    index ranges much longer than a warp."""
    from bramble_b200 import pack, synth
    m = common.MODES[mode]
    ann = synth.make_annotation(91, n_chrom=1, n_genes=6, hi_iso_genes=4, hi_iso_exons=14)
    assert max(np.bincount([int(t["gene"][1:].split(".")[0]) for t in ann.tx])) >= 50
    if m["long"]:
        recs = synth.make_long_reads(ann, None, 93, 600, params=dict(err=0.08))
    else:
        recs = synth.make_short_reads(ann, 93, 1500, multi_mean=2.0)
    ix = pack.build_index(ann)
    batches = pack.pack_batches(recs, ix, ann.chroms, long_reads=m["long"])
    wl = dict(index=ix, params=common.params_for(mode), batches=batches)
    n_rec = _check(wl, f"hot/{mode}")
    assert n_rec > 20 * len(recs) // 4


def test_contexts_in_parallel():
    """Several brb_ctx on one index, one host thread each (the end-to-end path of bench.py): bundle slices projected
    concurrently give exactly the rows of the whole bundle."""
    wl = common.make_workload("default", 43, n_names=4000)
    ix, p, b = wl["index"], wl["params"], wl["batches"][0]
    exp = oracle_lib.port_project(ix, p, b)
    G = b.desc.n_groups
    cuts = [G * i // 7 for i in range(8)]
    parts = [b.slice_groups(g0, g1) for g0, g1 in zip(cuts[:-1], cuts[1:])]
    with engine.Engine(ix) as eng:
        ctxs = [eng.context() for _ in range(3)]
        try:
            for _ in range(3):
                got = engine.project_pipelined(ctxs, p, parts, raw=False)
                assert sum(g["n_rec"] for g in got) == exp["n_rec"]
                for k in ("tid", "pos", "flags", "tx_strand"):
                    assert np.array_equal(np.concatenate([g[k] for g in got]), exp[k]), k
                a0 = np.concatenate([[0], np.cumsum([q.desc.n_aln for q in parts])])
                assert np.array_equal(np.concatenate([g["src_aln"] + a0[i] for i, g in enumerate(got)]), exp["src_aln"])
                cg = [c for g in got for c in g["cigars"]]
                assert all(np.array_equal(x, y) for x, y in zip(cg, exp["cigars"]))
        finally:
            for c in ctxs:
                c.close()


def test_rows_checksum_matches_oracle():
    """bench.py's full-size parity property: the order-independent row checksum of the engine's output (both row layouts,
    bundle slices through parallel contexts) equals the oracle's on the same bundle."""
    import ctypes as C
    wl = common.make_workload("default", 47, n_names=3000)
    ix, p, b = wl["index"], wl["params"], wl["batches"][0]
    o = oracle_lib.PortOut()
    assert oracle_lib.port().brb_port_project(C.byref(ix.desc), C.byref(p), C.byref(b.desc), C.byref(o)) == 0
    want = engine.rows_checksum(o)
    oracle_lib.port().brb_port_free(C.byref(o))
    assert want[1] > 1000
    G = b.desc.n_groups
    parts = [b.slice_groups(G * i // 5, G * (i + 1) // 5) for i in range(5)]
    with engine.Engine(ix) as eng:
        for ordered, packed in ((0, 0), (1, 0), (0, 1), (0, 2), (0, 3)):
            eng.set_option(capi.OPT_ORDERED_ROWS, ordered)
            eng.set_option(capi.OPT_PACKED_ROWS, packed)
            o = eng.project(p, b, raw=True)
            assert bool(o.packed_rows) == bool(packed) and bool(o.src_aln) != bool(packed)
            assert bool(o.aln_rec_cnt16) == (packed >= 2) and bool(o.aln_rec_off) != (packed >= 2) and bool(o.batch_rec_off) == (packed >= 2)
            assert o.packed_row_words == {0: 0, 1: 4, 2: 3, 3: 2}[packed]
            assert engine.rows_checksum(o, batch=b) == want
            if packed:                                         # bytes device -> host: 16 per row + 8 per alignment, or 12 + 2.1
                per_row, per_aln = {1: (16, 8), 2: (12, 2.125), 3: (8, 2.125)}[packed]
                assert abs(int(eng.stats().d2h_bytes) - (per_row * o.n_rec + per_aln * b.desc.n_aln + 4 * o.n_cigar_words)) < 256
        ctxs = [eng.context() for _ in range(2)]
        try:
            for packed in (0, 1, 2, 3):
                for c in ctxs:
                    c.set_option(capi.OPT_PACKED_ROWS, packed)
                res = engine.project_pipelined(ctxs, p, parts, checksum=True)
                assert (sum(r[4] for r in res) & 0xFFFFFFFFFFFFFFFF, sum(r[0] for r in res)) == want
        finally:
            for c in ctxs:
                c.close()


def test_packed_rows_oversized_cigar_word():
    """BRB_OPT_PACKED_ROWS keeps a one-word CIGAR inside the row record only below 2^27: a 9 Mb single-exon alignment
    (word 9000000M = 0x8954400 >= 2^27) must come back through the length-prefixed pool, next to ordinary inline rows."""
    from bramble_b200 import pack, synth
    ann = synth.Annotation(["chr1"], [20_000_000], [
        dict(id="G0.t0", gene="G0", chrom=0, strand="+", exons=[(1000, 12_000_000)], family=0),
        dict(id="G0.t1", gene="G0", chrom=0, strand="+", exons=[(1000, 5000), (6000, 12_000_000)], family=0),
        dict(id="G1.t0", gene="G1", chrom=0, strand="-", exons=[(13_000_000, 13_001_000), (13_002_000, 13_003_000)], family=1)])
    tags = [("NH", "C", 1), ("HI", "C", 1), ("AS", "C", 100), ("NM", "C", 0)]

    def rec(name, pos, cigar):
        return dict(name=name, flag=0, chrom=0, pos=pos, mapq=60, cigar=cigar, mchrom=-1, mpos=-1, isize=0, seq="", qual="", tags=tags)
    recs = [rec("a", 2000, [("M", 9_000_000)]),                           # inside t0's exon only ... and t1? no: it spans t1's intron
            rec("b", 7000, [("M", 9_000_000)]),                           # inside t0's exon and t1's second exon
            rec("c", 2000, [("M", 150)]),                                 # ordinary row, both transcripts
            rec("d", 13_000_900, [("M", 100), ("N", 999), ("M", 50)])]         # spliced, '-' transcript
    ix = pack.build_index(ann)
    p = common.params_for("default")
    b = pack.pack_batches(recs, ix, ann.chroms)[0]
    exp = oracle_lib.port_project(ix, p, b)
    assert exp["n_rec"] == 6 and any(len(c) == 1 and int(c[0]) >> 27 for c in exp["cigars"])
    with engine.Engine(ix) as eng:
        for packed in (0, 1):
            eng.set_option(capi.OPT_PACKED_ROWS, packed)
            oracle_lib.assert_rows_equal(exp, eng.project(p, b), f"oversized word packed={packed}")
        raw = eng.project(p, b, raw=True)
        assert raw.packed_rows and raw.n_cigar_words == 2 * sum(1 for c in exp["cigars"] if int(c[0]) >> 27)    # length word + CIGAR word each


def test_multiple_bundles_and_ragged():
    wl = common.make_workload("default", 41, n_names=3000, chunk_size=700)
    assert len(wl["batches"]) > 3
    _check(wl, "bundles")


def test_param_overrides():
    wl = common.make_workload("default", 51, n_names=2000)
    for kw in (dict(max_clip=0), dict(max_clip=25, max_junc_ins=3, max_junc_gap=3), dict(max_error_exon=35),
               dict(similarity_threshold=0.5, max_junc_ins=5, max_junc_gap=5)):
        wl["params"] = capi.resolve_params(**kw)
        _check(wl, f"override {kw}")
    wl = common.make_workload("lr_nofasta", 52, n_names=1000)
    for kw in (dict(lr=True, max_error_exon=0), dict(lr=True, similarity_threshold=0.2, max_junc_gap=100), dict(lr_hq=True)):
        wl["params"] = capi.resolve_params(**kw)
        _check(wl, f"override {kw}")


def test_empty_batch_and_no_match():
    wl = common.make_workload("default", 61, n_names=50)
    ix, p = wl["index"], wl["params"]
    from bramble_b200 import pack
    empty = pack.pack_batches([], ix, wl["ann"].chroms)
    assert empty == []
    b = pack.Batch().finish([], [], [], [0], [], [], [0], [], [], [0], None, None, [], [])
    with engine.Engine(ix) as eng:
        got = eng.project(p, b)
        assert got["n_rec"] == 0
        # alignments on an unannotated chromosome -> nothing projects
        recs = [dict(r, chrom=7) for r in wl["recs"] if not r["flag"] & 4][:20]
        for r in recs:
            r["mchrom"] = 7 if r["mchrom"] >= 0 else -1
        bs = pack.pack_batches(recs, ix, wl["ann"].chroms)
        got = eng.project(p, bs[0])
        assert got["n_rec"] == 0 and len(got["group_rec_off"]) == bs[0].desc.n_groups + 1


def test_lookup_kernel():
    wl = common.make_workload("lr_nofasta", 71, n_names=600)
    ix, p = wl["index"], wl["params"]
    refid, strand, qs, qe, status = [], [], [], [], []
    for b in wl["batches"]:
        A = b.arrays
        for a in range(b.desc.n_aln):
            e0, e1 = int(A["exon_off"][a]), int(A["exon_off"][a + 1])
            for j in range(e0, e1):
                n = e1 - e0
                st = 3 if n == 1 else (0 if j == e0 else (1 if j < e1 - 1 else 2))
                for sd in "+-":
                    refid.append(A["refid"][a]); strand.append(ord(sd)); qs.append(A["exon_st"][j]); qe.append(A["exon_en"][j]); status.append(st)
    refid = np.array(refid, np.int32); strand = np.array(strand, np.int8); qs = np.array(qs, np.uint32); qe = np.array(qe, np.uint32)
    status = np.array(status, np.uint8)
    cap = 64
    with engine.Engine(ix) as eng:
        hits, nh = eng.lookup(p, refid, strand, qs, qe, status, cap)
    eh = np.zeros((len(refid), cap), np.uint32); en = np.zeros(len(refid), np.uint32)
    import ctypes as C
    rc = oracle_lib.port().brb_port_lookup(C.byref(ix.desc), C.byref(p), len(refid), capi.ptr(refid), capi.ptr(strand), capi.ptr(qs),
                                           capi.ptr(qe), capi.ptr(status), cap, capi.ptr(eh.reshape(-1)), capi.ptr(en))
    assert rc == 0
    assert np.array_equal(nh, en)
    assert en.max() > 1
    for q in range(len(refid)):
        k = min(int(en[q]), cap)
        assert np.array_equal(hits[q, :k], eh[q, :k])


def _random_cigar_pairs(seed, n, max_real, max_ideal, zeros=False):
    rng = np.random.default_rng(seed)
    ops_r = [0, 1, 2, 3, 4, 5, 7, 8, 0, 0, 1, 2]
    ops_i = [0, 1, 2, 4, 10, 11, 12, 13, 0, 0, 2]
    reals, ideals = [], []
    for _ in range(n):
        nr, ni = int(rng.integers(0, max_real + 1)), int(rng.integers(0, max_ideal + 1))
        r = [(int(rng.integers(1, 40)) << 4) | int(rng.choice(ops_r)) for _ in range(nr)]
        i = [(int(rng.integers(0 if zeros else 1, 60)) << 4) | int(rng.choice(ops_i)) for _ in range(ni)]
        if nr and rng.random() < 0.4:                       # BAM-shaped ends: [H] [S] ... [S]
            r[0] = (r[0] >> 4 << 4) | (5 if rng.random() < 0.4 else 4)
            if nr > 1 and rng.random() < 0.5: r[1] = (r[1] >> 4 << 4) | 4
            if rng.random() < 0.5: r[-1] = (r[-1] >> 4 << 4) | 4
        if ni and rng.random() < 0.4:                       # clip-rescue segments lead with override ops
            for k in range(min(ni, 3)):
                i[k] = (i[k] >> 4 << 4) | int(rng.integers(10, 14))
        if zeros and nr and rng.random() < 0.1:             # a zero-length real op (not valid BAM): the cooperative code hands the pair to the serial one
            r[int(rng.integers(0, nr))] &= 0xf
        reals.append(np.array(r, np.uint32)); ideals.append(np.array(i, np.uint32))
    return reals, ideals


def _merge_on_gpu(eng, reals, ideals, warp):
    real_off = np.concatenate([[0], np.cumsum([len(x) for x in reals])]).astype(np.uint32)
    ideal_off = np.concatenate([[0], np.cumsum([len(x) for x in ideals])]).astype(np.uint32)
    real = np.concatenate(reals + [np.zeros(1, np.uint32)]).astype(np.uint32); ideal = np.concatenate(ideals + [np.zeros(1, np.uint32)]).astype(np.uint32)
    eng.set_option(capi.OPT_MERGE_WARP, int(warp))
    try:
        return eng.merge_cigars(real_off, real, ideal_off, ideal)
    finally:
        eng.set_option(capi.OPT_MERGE_WARP, 0)


@pytest.mark.parametrize("warp", [False, True])
def test_merge_kernel_golden_and_random(warp):
    """merge_cigars on the device -- one thread per pair (short-read rows) and the warp-cooperative code (long-read rows) -- against
    the oracle: the reference's 7 update_cigar vectors, 3000 short random pairs, 600 long ones (up to 700 real / 60 ideal ops)."""
    from test_oracle_golden import UPDATE_CIGAR_CASES, cig, port_merge
    wl = common.make_workload("default", 81, n_names=20)
    reals = [cig(r) for r, _, _ in UPDATE_CIGAR_CASES]
    ideals = [cig(i) for _, i, _ in UPDATE_CIGAR_CASES]
    for seed, n, mr, mi, z in ((5, 3000, 8, 8, False), (6, 400, 120, 20, True), (7, 200, 700, 60, False)):
        a, b = _random_cigar_pairs(seed, n, mr, mi, z)
        reals += a; ideals += b
    with engine.Engine(wl["index"]) as eng:
        got = _merge_on_gpu(eng, reals, ideals, warp)
    exp = port_merge(reals, ideals)
    for k, (g, e) in enumerate(zip(got, exp)):
        assert np.array_equal(g, e), (k, reals[k], ideals[k], g, e)
    for k, (_, _, want) in enumerate(UPDATE_CIGAR_CASES):
        assert np.array_equal(got[k], cig(want))


def test_ksw_kernel_golden_and_random():
    from test_oracle_golden import KSW_Q, KSW_T, codes, port_ksw
    wl = common.make_workload("default", 91, n_names=20)
    rq, rt = common.ksw_problems(17, 400)
    qs, ts = [codes(KSW_Q)] + rq, [codes(KSW_T)] + rt
    with engine.Engine(wl["index"]) as eng:
        got = eng.ksw_extend(qs, ts, cigar_cap=2048)
    exp = port_ksw(qs, ts, cap=2048)
    for k in ("max", "score", "max_q", "max_t", "zdropped", "n_cigar"):
        assert np.array_equal(got[k], exp[k]), k
    for i, (g, e) in enumerate(zip(got["cigars"], exp["cigars"])):
        assert np.array_equal(g, e), i
    assert (int(got["max"][0]), int(got["score"][0]), int(got["max_t"][0]), int(got["max_q"][0])) == (37, 11, 53, 54)
    assert exp["zdropped"].sum() > 0 and (exp["score"] > -0x40000000).sum() > 50


def test_pool_growth_replays(monkeypatch):
    """Pools start tiny: every stage overflows once, the host grows it and replays the kernel sequence."""
    monkeypatch.setenv("BRB_TEST_SMALL_POOLS", "1")
    wl = common.make_workload("default", 101, n_names=1500)
    _check(wl, "small pools")
    wl = common.make_workload("lr", 102, n_names=300)
    _check(wl, "small pools lr")


def test_bad_input_is_rejected():
    wl = common.make_workload("default", 111, n_names=200)
    ix, p = wl["index"], wl["params"]
    b = wl["batches"][0]
    with engine.Engine(ix) as eng:
        A = b.arrays
        m = A["mate_idx"]
        k = int(np.nonzero(m >= 0)[0][0])
        old = int(m[k]); m[k] = k                     # a read linked to itself
        with pytest.raises(capi.EngineError):
            eng.project(p, b)
        m[k] = old
        j = int(np.nonzero(np.diff(A["exon_off"].astype(np.int64)) >= 2)[0][0])
        e0 = int(A["exon_off"][j])
        s0, s1 = int(A["exon_st"][e0]), int(A["exon_st"][e0 + 1])
        A["exon_st"][e0 + 1] = s0                     # second exon starts inside the first
        with pytest.raises(capi.EngineError):
            eng.project(p, b)
        A["exon_st"][e0 + 1] = s1
        got = eng.project(p, b)                       # and the context is still usable
        oracle_lib.assert_rows_equal(oracle_lib.port_project(ix, p, b), got, "after bad input")


@pytest.mark.parametrize("stranded,multi_mean", [("fr", None), (None, None), (None, 5)])
def test_c2_shape_hot_genes(stranded, multi_mean):
    """Benchmark-shaped workload (GENCODE-like isoform counts up to 200 per gene, secondaries): the
    warp-flattened lookup and per-slot pairing on reads with hundreds of candidate transcripts."""
    from bramble_b200 import synth_fast as sf
    A = sf.make_annotation(7, n_chrom=3, n_genes=1500, iso_mean=6.0, iso_cap=200)
    ix = sf.build_index(A)
    B = sf.make_short_batch(A, ix, 8, 6000, stranded=stranded, multi_mean=multi_mean)
    p = capi.resolve_params()
    with engine.Engine(ix) as eng:
        exp = oracle_lib.port_project(ix, p, B)
        for generic in (0, 1):
            eng.set_option(capi.OPT_GENERIC_FRONT, generic)
            got = eng.project(p, B)
            oracle_lib.assert_rows_equal(exp, got, f"c2 shape generic={generic}")
            st = eng.stats()
            assert st.n_matches == exp["n_matches"] and got["n_rec"] > 10000
        assert exp["n_hits"] > 0


@pytest.mark.parametrize("kind", ["c5", "lr"])
def test_flush_tags_on_device_match_libstdcxx(kind):
    """BRB_OPT_FLUSH_TAGS: HI / NH / MAPQ and the primary choice per read name, computed on the device -- incl. the restated
    std::hash<std::string> -> std::mt19937_64 -> std::uniform_int_distribution tie-break -- against the oracle, which runs the
    reference's flush() arithmetic with the real libstdc++.  c5: multi-mappers, every multi-record name ties (scores are 0);
    lr: similarity scores decide, ties are rare."""
    if kind == "c5":
        from bramble_b200 import synth_fast as sf
        A = sf.make_annotation(5005, n_chrom=3, n_genes=600, paralog_families=60)
        ix = sf.build_index(A)
        b = sf.make_short_batch(A, ix, 77, 30000, stranded=None, multi_mean=5.0, paired_frac=0.6)
        names = [f"r{k:09d}" if k % 7 else f"read/{k}:a-rather-longer-name#{k * 977}" for k in range(int(b.desc.n_groups))]
        p = capi.resolve_params()
    else:
        wl = common.make_workload("lr", 61, n_names=4000, n_genes=120, n_chrom=2)
        ix, p, b = wl["index"], wl["params"], wl["batches"][0]
        go = b.arrays["group_off"]
        names = [b.names[int(go[g])] for g in range(int(b.desc.n_groups))]
    b.attach_names(names)
    exp = oracle_lib.port_project(ix, p, b)
    tags = oracle_lib.port_flush_tags(exp, b, bool(p.long_reads))
    with engine.Engine(ix) as eng:
        for ordered in (0, 1):
            eng.set_option(capi.OPT_ORDERED_ROWS, ordered)
            eng.set_option(capi.OPT_FLUSH_TAGS, 1)
            got = eng.project(p, b)
            oracle_lib.assert_rows_equal(exp, got, kind)
            for k in ("hi", "nh", "mapq", "primary"):
                assert np.array_equal(got[k], tags[k]), (kind, ordered, k, np.nonzero(got[k] != tags[k])[0][:5])
        eng.set_option(capi.OPT_PACKED_ROWS, 1); eng.set_option(capi.OPT_ORDERED_ROWS, 0)
        with pytest.raises(capi.EngineError):
            eng.project(p, b)
    assert int((tags["nh"] > 1).sum()) > 1000 and 0 < int(tags["primary"].sum()) < len(tags["primary"])


def _bgzf_blocks(buf):
    """[(payload, crc, isize)] of a BGZF byte string; checks the fixed header fields."""
    import struct
    out, p = [], 0
    while p < len(buf):
        assert buf[p:p + 4] == b"\x1f\x8b\x08\x04" and buf[p + 10:p + 16] == b"\x06\x00BC\x02\x00", p
        bsize = struct.unpack_from("<H", buf, p + 16)[0] + 1
        crc, isize = struct.unpack_from("<II", buf, p + bsize - 8)
        out.append((buf[p + 18:p + bsize - 8], crc, isize))
        p += bsize
    assert p == len(buf)
    return out


def test_bgzf_device_encoder_round_trips():
    """brb_bgzf_compress (LZ77 + fixed Huffman + CRC-32 on the device): every block inflates (zlib) to its 0xff00-byte share of the
    input, CRC32 and ISIZE are right, blocks stay within 64 KiB -- on incompressible, repetitive and BAM-shaped input."""
    import zlib
    from bramble_b200 import synth
    wl = common.make_workload("default", 131, n_names=400)
    rng = np.random.default_rng(9)
    bam = b"".join(synth._pack_record(r) for r in wl["recs"] if not r["flag"] & 4)
    cases = [b"", b"A", bytes(rng.integers(0, 256, size=200000, dtype=np.uint8)), b"ACGTTGCAAC" * 30000, bam, bam[:0xff00], bam[:0xff00 + 1],
             bytes(rng.integers(0, 4, size=150000, dtype=np.uint8)), (bam[:5000] * 40)[:0xff00 * 3 - 7],
             (bytes(rng.integers(0, 256, size=40000, dtype=np.uint8)) * 4)[:0xff00 * 2]]      # repeats 40 000 bytes apart: beyond deflate's 32 KiB window
    with engine.Engine(wl["index"]) as eng:
        for k, data in enumerate(cases):
            z = eng.bgzf_compress(data)
            blocks = _bgzf_blocks(z)
            assert len(blocks) == (len(data) + 0xff00 - 1) // 0xff00, k
            got = b""
            for payload, crc, isize in blocks:
                assert len(payload) + 26 <= 65536
                raw = zlib.decompress(payload, -15)
                assert len(raw) == isize and (zlib.crc32(raw) & 0xffffffff) == crc, k
                got += raw
            assert got == data, k
            if k in (3, 8):
                assert len(z) < len(data) // 4, (k, len(z), len(data))          # repeats are found (the BAM case has random bases and qualities)


# ---- brb_emit_bgzf: output records assembled on the device

def _aux_skip(a, s):
    """end of the aux value whose TYPE byte is at a[s], or None (htslib skip_aux)"""
    if s >= len(a):
        return None
    t = chr(a[s]); s += 1
    size = {"A": 1, "c": 1, "C": 1, "s": 2, "S": 2, "i": 4, "I": 4, "f": 4, "d": 8}
    if t in "ZH":
        e = a.find(b"\0", s)
        return None if e < 0 else e + 1
    if t == "B":
        if len(a) - s < 5:
            return None
        sz = size.get(chr(a[s]), 0)
        n = int.from_bytes(a[s + 1:s + 5], "little"); s += 5
        return None if (not sz or len(a) - s < sz * n) else s + sz * n
    sz = size.get(t, 0)
    return None if (not sz or len(a) - s < sz) else s + sz


def _aux_find(a, tag):
    s = 0
    while s is not None and len(a) - s >= 3:
        if a[s:s + 2] == tag:
            return s + 2
        s = _aux_skip(a, s + 2)
    return -1


def _aux_del(a, tag):
    o = _aux_find(a, tag)
    if o < 0:
        return a
    e = _aux_skip(a, o)
    return a if e is None else a[:o - 2] + a[e:]


def _py_record(body, lng, tid, pos, minus, paired, same_tx, primary, mapq, nh, hi, sim, clip, cigar, mate_tid, mate_pos):
    """write_to_bam of the reference for one row, tag edit by tag edit (src/core.cpp:96-212, src/bam.cpp:475-702): an independent
    restatement for this test (the product's host path is pinned on the reference binary by the CLI tests)."""
    import struct
    l_name, n_cig, flag, lq = body[8], struct.unpack_from("<H", body, 12)[0], struct.unpack_from("<H", body, 14)[0], struct.unpack_from("<i", body, 16)[0]
    p_seq = 32 + l_name + 4 * n_cig; seqb = (lq + 1) // 2
    seq, qual, aux = body[p_seq:p_seq + seqb], body[p_seq + seqb:p_seq + seqb + lq], bytes(body[p_seq + seqb + lq:])
    aux = _aux_del(aux, b"NH") + b"NHi" + struct.pack("<i", nh)
    aux = _aux_del(aux, b"ts" if lng else b"XS")
    if lng:
        o = _aux_find(aux, b"AS"); v = 0
        if o >= 0:
            t = chr(aux[o]); fmt = {"c": "<b", "C": "<B", "s": "<h", "S": "<H", "i": "<i", "I": "<I"}.get(t)
            v = struct.unpack_from(fmt, aux, o + 1)[0] if fmt else 0
            aux = _aux_del(aux, b"AS")
        aux += b"ASi" + struct.pack("<i", int((float(v) + float(clip)) * sim))
    aux = _aux_del(aux, b"HI") + b"HIi" + struct.pack("<i", hi)
    flag = (flag & ~0x100) if primary else (flag | 0x100)
    if minus:
        if lq > 0:
            comp = {1: 8, 8: 1, 2: 4, 4: 2}
            nib = [(seq[j >> 1] >> (0 if j & 1 else 4)) & 15 for j in range(lq)]
            rc = [comp.get(n, 15) for n in reversed(nib)] + [0]
            seq = bytes((rc[2 * k] << 4) | rc[2 * k + 1] for k in range(seqb))
            if qual[0] != 0xff:
                qual = qual[::-1]
        flag ^= 16
    if not paired:
        flag &= ~(1 | 2 | 32); mtid, mpos, isize = -1, -1, 0
    else:
        flag |= 1
        if minus:
            flag |= 32
        if same_tx:
            mtid, mpos = tid, mate_pos; flag |= 2
            isize = (mate_pos + lq) - pos if pos <= mate_pos else -((pos + lq) - mate_pos)
        else:
            mtid, mpos, isize = mate_tid, mate_pos, 0; flag &= ~2
    head = bytearray(body[:32 + l_name])
    struct.pack_into("<ii", head, 0, tid, pos); head[9] = mapq & 0xff
    struct.pack_into("<HH", head, 12, len(cigar), flag); struct.pack_into("<iii", head, 20, mtid, mpos, isize)
    out = bytes(head) + struct.pack("<%dI" % len(cigar), *[int(c) for c in cigar]) + seq + qual + aux
    return struct.pack("<i", len(out)) + out


@pytest.mark.parametrize("mode", ["default", "lr"])
def test_emit_records_on_device_match_restatement(mode):
    """brb_emit_bgzf: BAM records built on the device from the resident rows + the input records, then BGZF: inflated, they must be the
    records a tag-by-tag Python restatement of write_to_bam builds from the downloaded rows -- random aux tag sets (NH / HI / AS / XS / ts
    in every integer type, duplicates, B / Z / H / f tags, a malformed tail), odd read lengths, missing qualities, arbitrary unit order
    and tag values.  Also: the rows moved to another ctx first (brb_ctx_move_result), and the error paths."""
    import zlib
    from bramble_b200 import synth
    wl = common.make_workload(mode, 211, n_names=700, n_genes=40)
    lng = wl["mode"]["long"]
    rng = np.random.default_rng(5)
    b = wl["batches"][0]
    pool = [("NH", "C", 3), ("NH", "i", 70000), ("HI", "s", -2), ("HI", "C", 9), ("AS", "c", -7), ("AS", "S", 40000), ("AS", "i", 250), ("AS", "I", 12), ("NM", "C", 1),
            ("MD", "Z", "50A49"), ("XH", "H", "0AFF"), ("XB", "B", ("S", [1, 2, 65535])), ("XE", "B", ("i", [])), ("XF", "f", 0.5), ("ZZ", "Z", "")]
    bodies = []
    for a, ri in enumerate(b.src):
        r = dict(wl["recs"][int(ri)])
        strand_tags = [t for t in r.get("tags", []) if t[0] in ("XS", "ts")]            # (they decide the strand: kept as they are, in front)
        extra = [pool[int(k)] for k in rng.permutation(len(pool))[:int(rng.integers(0, 9))]]
        r["tags"] = strand_tags + extra + ([("ts", "A", "+")] if lng and a % 5 == 0 else []) + ([("XS", "A", "-")] if a % 7 == 0 and strand_tags else [])
        if a % 11 == 0:
            r["qual"] = None
        body = synth._pack_record(r)[4:]
        if a % 13 == 0:
            body += b"XXZunterminated"                                                   # a tag no walk can skip: everything behind it is out of reach
        bodies.append(body)
    rec_len = np.array([len(x) for x in bodies], np.uint32)
    rec_off = (np.cumsum(rec_len) - rec_len).astype(np.uint64)
    raw = b"".join(bodies)
    p = wl["params"]
    with engine.Engine(wl["index"]) as eng:
        out = eng.project(p, b, raw=True)                          # rows in the device's own order: that is what a unit's row index means
        n = int(out.n_rec)
        o = engine._columns(out, copy=True)
        co = o["cigar_off"]
        assert n > 500
        # units in a shuffled order with arbitrary tag values
        flags = o["flags"]; firsts = [r for r in range(n) if flags[r] & 1]
        order = rng.permutation(len(firsts))
        units = np.array([[firsts[k], int(rng.integers(1, 50)), int(rng.integers(1, 90)), int(rng.integers(0, 256)) | (int(rng.integers(0, 2)) << 8)] for k in order], np.uint32)
        want = b""
        for row, hi, nh, mq in units:
            paired, same = bool(flags[row] & 2), bool(flags[row] & 4)
            for e in range(2 if paired else 1):
                r, other = int(row) + e, int(row) + (1 - e)
                a = int(o["src_aln"][r])
                want += _py_record(bodies[a], lng, int(o["tid"][r]), int(o["pos"][r]), o["tx_strand"][r] == ord("-"), paired, same, bool(mq >> 8), int(mq) & 255, int(nh), int(hi) + e,
                                   float(o["sim_score"][r]), int(o["clip_score"][r]), o["cigar"][int(co[r]):int(co[r + 1])], int(o["tid"][other]) if paired else -1, int(o["pos"][other]) if paired else -1)

        def inflate(z):
            return b"".join(zlib.decompress(pl, -15) for pl, _, _ in _bgzf_blocks(z))
        z, rawn = eng.emit_bgzf(p, raw, rec_off, rec_len, units)
        got = inflate(z)
        assert rawn == len(want) and got == want, (mode, rawn, len(want), next((i for i, (x, y) in enumerate(zip(got, want)) if x != y), None))
        # the same from a second context that only received the rows
        holder = engine.Context(eng)                               # (the rows of the projection above: a new one may finish its batches in another order)
        eng.move_result(holder)
        z2, _ = eng.emit_bgzf(p, raw, rec_off, rec_len, units, ctx=holder)
        assert inflate(z2) == want
        with pytest.raises(capi.EngineError):                      # ... and the source holds no result any more
            eng.emit_bgzf(p, raw, rec_off, rec_len, units)
        # error paths: a unit naming a row that does not exist; a record whose fixed fields overrun its length
        out = eng.project(p, b, raw=True)
        o = engine._columns(out, copy=True)
        units = np.array([[r, 1, 1, 255] for r in range(n) if o["flags"][r] & 1], np.uint32)
        bad = units.copy(); bad[3, 0] = n + 5
        with pytest.raises(capi.EngineError):
            eng.emit_bgzf(p, raw, rec_off, rec_len, bad)
        short = rec_len.copy(); short[int(o["src_aln"][int(units[0, 0])])] = 40
        with pytest.raises(capi.EngineError):
            eng.emit_bgzf(p, raw, rec_off, short, units)
        z3, _ = eng.emit_bgzf(p, raw, rec_off, rec_len, units[:0])
        assert z3 == b""
        holder.close()


def test_bgzf_device_decoder_matches_zlib():
    """brb_bgzf_inflate (Huffman decode + LZ77 copies + CRC-32 check on the device, one warp per block): BGZF blocks written by zlib at
    levels 0 / 1 / 6 / 9 (stored, fixed and dynamic blocks), by the device encoder (fixed codes), with several deflate blocks inside one
    BGZF block, with long overlapping matches, plus the empty EOF block -- must come back as the original bytes; a flipped payload byte
    and a wrong CRC must be refused."""
    import struct
    import zlib
    from bramble_b200 import synth
    wl = common.make_workload("default", 133, n_names=500)
    rng = np.random.default_rng(11)
    bam = b"".join(synth._pack_record(r) for r in wl["recs"] if not r["flag"] & 4)

    def blocks(data, level, step=0xff00):
        return b"".join(synth._bgzf_block(data[i:i + step], level) for i in range(0, len(data), step))

    def multi(data, level):                                  # several deflate blocks (sync flushes add empty stored blocks) in ONE BGZF block
        co = zlib.compressobj(level, zlib.DEFLATED, -15)
        k = len(data) // 3
        c = co.compress(data[:k]) + co.flush(zlib.Z_SYNC_FLUSH) + co.compress(data[k:2 * k]) + co.flush(zlib.Z_FULL_FLUSH) + co.compress(data[2 * k:]) + co.flush()
        return (b"\x1f\x8b\x08\x04\0\0\0\0\0\xff\x06\0BC\x02\0" + struct.pack("<H", len(c) + 25) + c + struct.pack("<II", zlib.crc32(data) & 0xffffffff, len(data)))

    noise = bytes(rng.integers(0, 256, size=150000, dtype=np.uint8))
    runs = b"A" * 70000 + b"ACGT" * 20000 + bytes(rng.integers(0, 4, size=60000, dtype=np.uint8)) + b"\0" * 65000
    with engine.Engine(wl["index"]) as eng:
        cases = []
        for level in (0, 1, 6, 9):
            cases.append((f"bam level {level}", bam, blocks(bam, level)))
            cases.append((f"runs level {level}", runs, blocks(runs, level)))
        cases.append(("noise level 6", noise, blocks(noise, 6)))
        cases.append(("small blocks", bam[:50000], blocks(bam[:50000], 6, step=777)))
        cases.append(("multi", bam[:60000], multi(bam[:60000], 6) + multi(runs[:65000], 1)))
        cases[-1] = ("multi", bam[:60000] + runs[:65000], cases[-1][2])
        cases.append(("device encoder", bam, eng.bgzf_compress(bam)))
        cases.append(("with EOF block", bam[:1000], blocks(bam[:1000], 6) + synth._BGZF_EOF))
        cases.append(("only EOF", b"", synth._BGZF_EOF))
        cases.append(("nothing", b"", b""))
        for name, want, z in cases:
            got = eng.bgzf_inflate(z)
            assert got == want, (name, len(got), len(want), next((i for i, (x, y) in enumerate(zip(got, want)) if x != y), None))
        z = bytearray(blocks(bam, 6))
        z[len(z) // 2] ^= 0x55                                   # somewhere inside a payload
        with pytest.raises(capi.EngineError):
            eng.bgzf_inflate(bytes(z))
        z = bytearray(blocks(bam[:40000], 6))
        z[-8] ^= 1                                              # the footer's CRC-32
        with pytest.raises(capi.EngineError, match="CRC"):
            eng.bgzf_inflate(bytes(z))
        with pytest.raises(capi.EngineError):
            eng.bgzf_inflate(blocks(bam[:40000], 6)[:-5])       # truncated
        assert eng.bgzf_inflate(blocks(bam, 1)) == bam          # the context is fine after the refusals


# ---- compact input columns (brb_batch_in::pos0 / cigar_cnt16 / mate_delta16 / refid16)

def test_read_exons_derived_on_device_match_ingest():
    """pos0 + CIGAR -> read exons on the device (k_expand.cuh) against the Python restatement of the reference's ingest (pack.read_exons, which
    the CLI tests pin on the reference binary): every CIGAR shape incl. the quirks -- leading N, I between two Ns (no exon closed), D, clips,
    =/X, N at the end -- and random ones."""
    from bramble_b200 import pack
    wl = common.make_workload("default", 141, n_names=20)
    rng = np.random.default_rng(3)
    M, I, D, N, S, H, EQ, X = 0, 1, 2, 3, 4, 5, 7, 8
    shapes = [[(M, 100)], [(S, 5), (M, 50), (N, 200), (M, 45)], [(N, 30), (M, 60)], [(M, 40), (N, 100), (I, 3), (N, 50), (M, 57)],
              [(M, 40), (I, 3), (N, 100), (M, 57)], [(H, 10), (S, 2), (EQ, 30), (X, 1), (EQ, 20), (D, 4), (M, 10), (N, 1000), (M, 30), (S, 7)],
              [(M, 10), (N, 20)], [(M, 10), (N, 20), (N, 30), (M, 5)], [(I, 5), (M, 20)], [(M, 20), (D, 5), (N, 7), (D, 2), (M, 9)], [(M, 1)]]
    for _ in range(400):
        k = int(rng.integers(1, 12))
        shapes.append([(int(rng.choice([M, M, M, I, D, N, N, S, EQ, X])), int(rng.integers(1, 300))) for _ in range(k)])
    cigs, pos, exp_off, exp_st, exp_en = [], [], [0], [], []
    for sh in shapes:
        words = [(l << 4) | o for o, l in sh]
        p0 = int(rng.integers(0, 1 << 27))
        ex = pack.read_exons(p0, words)
        cigs.append(np.array(words, np.uint32)); pos.append(p0)
        exp_st += [s for s, _ in ex]; exp_en += [e for _, e in ex]; exp_off.append(len(exp_st))
    n = len(cigs)
    assert n > 300
    b = pack.Batch()
    A = b.arrays
    A["group_off"] = np.arange(n + 1, dtype=np.uint32); A["strand"] = np.full(n, ord("."), np.int8); A["cigar"] = np.concatenate(cigs)
    A["cigar_cnt16"] = np.array([len(c) for c in cigs], np.uint16); A["mate_delta16"] = np.zeros(n, np.int16); A["refid16"] = np.zeros(n, np.int16)
    A["pos0"] = np.array(pos, np.int32)
    b.desc.n_aln = b.desc.n_groups = n; b.desc.n_cigar_total = int(len(A["cigar"]))
    for k, a in A.items():
        setattr(b.desc, k, capi.ptr(a))
    with engine.Engine(wl["index"]) as eng:
        eng.upload(b)
        off, st, en = eng.read_exons(len(exp_st) + 10)
        assert np.array_equal(off, np.array(exp_off, np.uint32)) and np.array_equal(st, np.array(exp_st, np.uint32)) and np.array_equal(en, np.array(exp_en, np.uint32))
        # no aligned base at position 0 (the last exon end stays 0): the reference's "invalid CIGAR" error
        A["cigar"][0] = (50 << 4) | S; A["cigar_cnt16"][0] = 1; A["pos0"][0] = 0
        A["cigar"] = np.concatenate([A["cigar"][:1], np.concatenate(cigs[1:])]); b.desc.cigar = capi.ptr(A["cigar"]); b.desc.n_cigar_total = int(len(A["cigar"]))
        with pytest.raises(capi.EngineError, match="invalid CIGAR"):
            eng.upload(b)
        b.desc.n_cigar_total += 1
        with pytest.raises(capi.EngineError, match="n_cigar_total"):
            eng.upload(b)


@pytest.mark.parametrize("mode", ["default", "fr", "lr", "lrhq"])
def test_compact_input_columns_give_the_same_rows(mode):
    """A bundle uploaded with the compact columns (16-bit CIGAR counts, mate deltas, refids; position instead of the exon arrays) projects
    to the rows of the plain columns = the oracle's; fewer bytes go up."""
    wl = common.make_workload(mode, 143, n_names=1200, n_genes=40)
    ix, p = wl["index"], wl["params"]
    with engine.Engine(ix) as eng:
        for b in wl["batches"]:
            exp = oracle_lib.port_project(ix, p, b)
            full = eng.project(p, b); h_full = int(eng.stats().h2d_bytes)
            oracle_lib.assert_rows_equal(exp, full, f"{mode} plain")
            for exons in (True, False):
                got = eng.project(p, b.compact(exons=exons))
                oracle_lib.assert_rows_equal(exp, got, f"{mode} compact exons={exons}")
                assert int(eng.stats().h2d_bytes) < h_full
            eng.set_option(capi.OPT_PACKED_ROWS, 3 if not wl["mode"]["long"] else 1)
            oracle_lib.assert_rows_equal(exp, eng.project(p, b.compact()), f"{mode} compact in, packed out")
            eng.set_option(capi.OPT_PACKED_ROWS, 0)


def test_compact_input_edge_cases():
    """Empty bundle and single-alignment bundles through the compact columns; a compact upload followed by a plain one on the same context."""
    from bramble_b200 import pack
    wl = common.make_workload("default", 145, n_names=300, n_genes=20)
    ix, p, b = wl["index"], wl["params"], wl["batches"][0]
    with engine.Engine(ix) as eng:
        empty = b.slice_groups(0, 0)
        assert eng.project(p, empty.compact())["n_rec"] == 0
        exp = oracle_lib.port_project(ix, p, b)
        for g in range(0, int(b.desc.n_groups), max(1, int(b.desc.n_groups) // 12)):
            one = b.slice_groups(g, g + 1)
            oracle_lib.assert_rows_equal(oracle_lib.port_project(ix, p, one), eng.project(p, one.compact()), f"group {g}")
        oracle_lib.assert_rows_equal(exp, eng.project(p, b.compact()), "compact")
        oracle_lib.assert_rows_equal(exp, eng.project(p, b), "plain after compact")
        oracle_lib.assert_rows_equal(exp, eng.project(p, b.compact(exons=False)), "compact again")
