"""Thin object wrapper over the C ABI: index + context lifetime, project(), stats().
Mirrors the reference seam `convert_reads(reads, g2t, evaluator, ...)` (src/core.cpp:220): the index is
the g2t tree, a Batch is one bundle, the returned rows are the BamInfo ends in emit order."""
import ctypes as C

import numpy as np

from . import capi


def _unit_blocks(o, n_aln):
    """(first row, rows) of the unit block owned by every alignment: aln_rec_off / aln_rec_cnt, or -- BRB_OPT_PACKED_ROWS = 2 --
    rebuilt from one row offset per batch of 32 alignments and the 16-bit unit sizes."""
    if getattr(o, "aln_rec_cnt16", None):
        cnt = np.ctypeslib.as_array(o.aln_rec_cnt16, shape=(n_aln,)).astype(np.int64)
        nb = (n_aln + 31) // 32
        boff = np.ctypeslib.as_array(o.batch_rec_off, shape=(nb,)).astype(np.int64)
        pad = np.zeros(nb * 32, np.int64); pad[:n_aln] = cnt
        within = np.cumsum(pad.reshape(nb, 32), axis=1) - pad.reshape(nb, 32)
        return (boff[:, None] + within).reshape(-1)[:n_aln], cnt
    if n_aln == 0:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    return (np.ctypeslib.as_array(o.aln_rec_off, shape=(n_aln,)).astype(np.int64), np.ctypeslib.as_array(o.aln_rec_cnt, shape=(n_aln,)).astype(np.int64))


def _columns(o, copy=False, batch=None):
    """Row columns of a raw brb_batch_out in MEMORY order, whichever layout it came in: the structure-of-arrays columns,
    or BRB_OPT_PACKED_ROWS records (include/bramble_b200.h) unpacked into the same columns (a dense CIGAR CSR rebuilt
    from the in-record words and the length-prefixed pool).  The compact layout (option value 2) leaves the source alignment
    out: it follows from the unit blocks and the batch's mate links, so `batch` is needed."""
    n, W = int(o.n_rec), int(o.n_cigar_words)

    def arr(p, cnt, dt=None):
        if cnt == 0 or not p:          # NULL score columns: all zero (see include/bramble_b200.h)
            return np.zeros(cnt, dtype=dt or np.uint32)
        a = np.ctypeslib.as_array(p, shape=(cnt,))
        return a.astype(dt or a.dtype, copy=True) if (copy or dt) else a
    d = dict(sim_score=arr(o.sim_score, n, np.float64), clip_score=arr(o.clip_score, n, np.int32))
    if not getattr(o, "packed_rows", None):      # (the oracle's result struct has the plain columns only)
        d.update(src_aln=arr(o.src_aln, n), tid=arr(o.tid, n), pos=arr(o.pos, n), tx_strand=arr(o.tx_strand, n, np.int8),
                 flags=arr(o.flags, n, np.uint8), cigar_off=arr(o.cigar_off, n + 1).astype(np.int64), cigar=arr(o.cigar, W))
        return d
    compact = bool(getattr(o, "aln_rec_cnt16", None))
    if compact:
        assert batch is not None, "compact rows: the batch is needed to name the source alignments"
        n_aln = int(batch.desc.n_aln)
        aoff, acnt = _unit_blocks(o, n_aln)
        owner = np.zeros(n, np.int64)
        owner[(np.repeat(aoff, acnt) + (np.arange(int(acnt.sum())) - np.repeat(np.cumsum(acnt) - acnt, acnt)))] = np.repeat(np.arange(n_aln), acnt)
        mates = getattr(batch, "arrays_full", batch.arrays)["mate_idx"].astype(np.int64)
        if int(getattr(o, "packed_row_words", 3)) == 2:                 # option value 3: one 64-bit word per row (include/bramble_b200.h, BRB_ROW8_*)
            v = arr(o.packed_rows, 2 * n).view(np.uint64)
            tid8 = (v & np.uint64(0x3FFFFF)).astype(np.uint32); pos8 = ((v >> np.uint64(22)) & np.uint64(0xFFFFFF)).astype(np.uint32)
            fl8 = ((v >> np.uint64(46)) & np.uint64(7)).astype(np.uint8); minus8 = ((v >> np.uint64(49)) & np.uint64(1)).astype(bool)
            inl8 = ((v >> np.uint64(50)) & np.uint64(1)).astype(bool); n8 = (v >> np.uint64(51)).astype(np.int64)
            pool = arr(o.cigar, W)
            clen = np.where(inl8, 1, n8)
            coff = np.concatenate([[0], np.cumsum(clen)])
            words = np.zeros(int(coff[-1]), dtype=np.uint32)
            words[coff[:-1][inl8]] = (n8[inl8].astype(np.uint32) << np.uint32(4))               # (len << 4) | M
            lens = n8[~inl8]
            if lens.size:                                               # pool entries in row order: entry k starts behind the k earlier ones (and their length words)
                start = np.cumsum(lens + 1) - (lens + 1)
                assert np.array_equal(pool[start], lens.astype(np.uint32)), "8-byte rows: pool entries are not in row order"
                within = np.arange(int(lens.sum()), dtype=np.int64) - np.repeat(np.cumsum(lens) - lens, lens)
                words[np.repeat(coff[:-1][~inl8], lens) + within] = pool[np.repeat(start + 1, lens) + within]
            src = np.where((fl8 & 1) == 1, owner, mates[owner]).astype(np.uint32)
            d.update(src_aln=src, tid=tid8, pos=pos8, tx_strand=np.where(minus8, ord("-"), ord("+")).astype(np.int8), flags=fl8, cigar_off=coff, cigar=words)
            return d
        p3 = arr(o.packed_rows, 3 * n).reshape(n, 3)
        first = ((p3[:, 2] >> 29) & 1) == 1
        src = np.where(first, owner, mates[owner]).astype(np.uint32)
        pr = np.concatenate([src[:, None], p3], axis=1)
    else:
        pr = arr(o.packed_rows, 4 * n).reshape(n, 4)
    w = pr[:, 3]
    inline, val = ((w >> 27) & 1) == 1, (w & 0x07FFFFFF).astype(np.int64)
    pool = arr(o.cigar, W)
    clen = np.ones(n, dtype=np.int64)
    clen[~inline] = pool[val[~inline]]
    coff = np.concatenate([[0], np.cumsum(clen)])
    words = np.zeros(int(coff[-1]), dtype=np.uint32)
    words[coff[:-1][inline]] = val[inline].astype(np.uint32)
    lens = clen[~inline]
    if lens.size:
        within = np.arange(int(lens.sum()), dtype=np.int64) - np.repeat(np.cumsum(lens) - lens, lens)
        words[np.repeat(coff[:-1][~inline], lens) + within] = pool[np.repeat(val[~inline] + 1, lens) + within]
    d.update(src_aln=pr[:, 0].copy(), tid=pr[:, 1].copy(), pos=pr[:, 2].copy(),
             tx_strand=np.where((w >> 28) & 1, ord("-"), ord("+")).astype(np.int8), flags=(w >> 29).astype(np.uint8),
             cigar_off=coff, cigar=words)
    return d


def _rows(o, batch):
    """Rows of a brb_batch_out in the reference's emit order: the unit blocks (aln_rec_off / aln_rec_cnt) concatenated in
    alignment order; group_rec_off follows from the block sizes."""
    n = int(o.n_rec)
    n_aln, n_groups = int(batch.desc.n_aln), int(batch.desc.n_groups)

    def arr(p, cnt, dt):
        if cnt == 0 or not p:
            return np.zeros(cnt, dtype=dt)
        return np.ctypeslib.as_array(p, shape=(cnt,)).astype(dt, copy=True)
    aoff, acnt = _unit_blocks(o, n_aln)
    assert int(acnt.sum()) == n, "unit blocks do not cover the rows"
    start = np.cumsum(acnt) - acnt
    order = np.repeat(aoff - start, acnt) + np.arange(n, dtype=np.int64)        # row index in memory of the k-th row in emit order
    go = batch.arrays["group_off"].astype(np.int64)[:n_groups + 1]
    cum = np.concatenate([[0], np.cumsum(acnt)])
    col = _columns(o, copy=True, batch=batch)
    d = dict(n_rec=n, group_rec_off=cum[go].astype(np.uint32) if n_aln else np.zeros(n_groups + 1, np.uint32))
    for k in ("src_aln", "tid", "pos", "tx_strand", "flags", "sim_score", "clip_score"):
        d[k] = col[k][order]
    if getattr(o, "hi", None):                                                   # BRB_OPT_FLUSH_TAGS
        for k, dt in (("hi", np.int32), ("nh", np.int32), ("mapq", np.uint8), ("primary", np.uint8)):
            d[k] = arr(getattr(o, k), n, dt)[order]
    if o.group_rec_off:                                                          # BRB_OPT_ORDERED_ROWS: already in emit order
        assert np.array_equal(order, np.arange(n)) and np.array_equal(arr(o.group_rec_off, n_groups + 1, np.uint32), d["group_rec_off"])
    co, pool = col["cigar_off"], col["cigar"]
    d["cigar_len"] = np.diff(co).astype(np.uint32)[order]
    d["cigars"] = [pool[int(co[r]):int(co[r + 1])].copy() for r in order]
    d["cigar_off"] = (np.cumsum(d["cigar_len"]) - d["cigar_len"]).astype(np.uint32)
    return d


class Context:
    """An extra brb_ctx on an Engine's index: own CUDA stream, scratch pools and pinned staging.  The C ABI wants
    one context per host thread; several of them overlap the H2D copy of one bundle with the kernels of the
    next and the D2H copy of the previous one (the reference runs its bundles on `-p` worker threads the same
    way, src/threads.cpp:100-104)."""

    def __init__(self, eng):
        self.lib = eng.lib
        self._ctx = C.c_void_p()
        capi.check(self.lib.brb_ctx_create(eng._ix, C.byref(self._ctx)), "brb_ctx_create")

    def project(self, params, batch, raw=False):
        out = capi.BatchOut()
        capi.check(self.lib.brb_project(self._ctx, C.byref(params), C.byref(batch.desc), C.byref(out)), "brb_project")
        return out if raw else _rows(out, batch)

    def stats(self):
        s = capi.Stats()
        capi.check(self.lib.brb_ctx_stats(self._ctx, C.byref(s)), "brb_ctx_stats")
        return s

    def set_option(self, option, value):
        capi.check(self.lib.brb_ctx_set_option(self._ctx, int(option), int(value)), "brb_ctx_set_option")

    def close(self):
        if self._ctx:
            self.lib.brb_ctx_destroy(self._ctx)
            self._ctx = C.c_void_p()


def rows_checksum(o, src_offset=0, batch=None):
    """Order-independent 64-bit checksum of the rows of a raw brb_batch_out (sum of a per-row mix of source alignment,
    tid, pos, strand, flags, similarity score bits, clip score, CIGAR length and CIGAR words): equal for any layout of
    the same row set."""
    n = int(o.n_rec)
    if n == 0:
        return 0, 0
    c = _columns(o, batch=batch)
    M = np.uint64(0x9E3779B97F4A7C15)
    with np.errstate(over="ignore"):
        co = c["cigar_off"].astype(np.uint64)
        pool = c["cigar"].astype(np.uint64)
        wsum = np.concatenate([[np.uint64(0)], np.cumsum(pool * M + (pool >> np.uint64(7)), dtype=np.uint64)])
        h = (c["src_aln"].astype(np.uint64) + np.uint64(src_offset)) * M
        h ^= c["tid"].astype(np.uint64) * np.uint64(0xC2B2AE3D27D4EB4F)
        h = h * M + c["pos"].astype(np.uint64)
        h = h * M + (c["tx_strand"].astype(np.uint8).astype(np.uint64) << np.uint64(8)) + c["flags"].astype(np.uint64)
        h = h * M + np.ascontiguousarray(c["sim_score"], dtype=np.float64).view(np.uint64)
        h = h * M + c["clip_score"].astype(np.int64).astype(np.uint64)
        h = h * M + (co[1:] - co[:-1])
        h = h * M + (wsum[co[1:].astype(np.int64)] - wsum[co[:-1].astype(np.int64)])
        return int(h.sum(dtype=np.uint64)), n


def project_pipelined(contexts, params, bundles, raw=True, checksum=False):
    """Bundles through `contexts` (one host thread each, ctypes releases the GIL): results in bundle order.
    A context's result arrays are only valid until its next call, so raw results have to be consumed by
    the caller before that; with raw=True they are reduced here to (n_rec, n_cigar_words, h2d_bytes, d2h_bytes, rows_checksum or 0)."""
    from concurrent.futures import ThreadPoolExecutor
    import queue
    free = queue.SimpleQueue()
    for c in contexts:
        free.put(c)

    def run(b):
        c = free.get()
        try:
            o = c.project(params, b, raw=raw)
            if not raw:
                return o
            st = c.stats()
            cs = rows_checksum(o, b.src_offset, batch=b)[0] if checksum else 0
            return (int(o.n_rec), int(o.n_cigar_words), int(st.h2d_bytes), int(st.d2h_bytes), cs)
        finally:
            free.put(c)
    with ThreadPoolExecutor(max_workers=len(contexts)) as ex:
        return list(ex.map(run, bundles))


class Engine:
    def __init__(self, index, device=0):
        self.lib = capi.lib()
        if self.lib.brb_device_count() <= 0:
            raise capi.EngineError("no CUDA device: bramble_b200 has no CPU path")
        self.index = index
        self._ix = C.c_void_p()
        self._ctx = C.c_void_p()
        capi.check(self.lib.brb_index_create(device, C.byref(index.desc), C.byref(self._ix)), "brb_index_create")
        capi.check(self.lib.brb_ctx_create(self._ix, C.byref(self._ctx)), "brb_ctx_create")

    def close(self):
        if self._ctx:
            self.lib.brb_ctx_destroy(self._ctx)
            self._ctx = C.c_void_p()
        if self._ix:
            self.lib.brb_index_destroy(self._ix)
            self._ix = C.c_void_p()

    def context(self):
        return Context(self)

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def project(self, params, batch, raw=False):
        out = capi.BatchOut()
        capi.check(self.lib.brb_project(self._ctx, C.byref(params), C.byref(batch.desc), C.byref(out)), "brb_project")
        return out if raw else _rows(out, batch)

    def upload(self, batch):
        capi.check(self.lib.brb_batch_upload(self._ctx, C.byref(batch.desc)), "brb_batch_upload")
        self._n_aln_resident = int(batch.desc.n_aln)

    def read_exons(self, cap):
        """(exon_off, exon_st, exon_en) of the resident batch as the engine holds them."""
        n_aln = int(self._n_aln_resident)
        off = np.zeros(n_aln + 1, np.uint32); st = np.zeros(max(1, cap), np.uint32); en = np.zeros(max(1, cap), np.uint32); n = C.c_uint64()
        capi.check(self.lib.brb_batch_read_exons(self._ctx, capi.ptr(off), capi.ptr(st), capi.ptr(en), cap, C.byref(n)), "brb_batch_read_exons")
        return off, st[:n.value], en[:n.value]

    def run_resident(self, params):
        capi.check(self.lib.brb_project_resident(self._ctx, C.byref(params)), "brb_project_resident")

    def download(self, batch, raw=False):
        out = capi.BatchOut()
        capi.check(self.lib.brb_batch_download(self._ctx, C.byref(out)), "brb_batch_download")
        return out if raw else _rows(out, batch)

    def set_option(self, option, value):
        capi.check(self.lib.brb_ctx_set_option(self._ctx, option, int(value)), "brb_ctx_set_option")

    def timer_start(self):
        capi.check(self.lib.brb_ctx_timer_start(self._ctx), "brb_ctx_timer_start")

    def timer_stop(self):
        ms = C.c_float()
        capi.check(self.lib.brb_ctx_timer_stop(self._ctx, C.byref(ms)), "brb_ctx_timer_stop")
        return float(ms.value)

    def pinned_copy(self, batch):
        """Copy of `batch` whose columns live in pinned host memory (brb_host_alloc)."""
        import copy
        nb = copy.copy(batch)
        nb.desc = capi.BatchIn()
        nb.desc.n_aln, nb.desc.n_groups = batch.desc.n_aln, batch.desc.n_groups
        nb.desc.n_cigar_total = batch.desc.n_cigar_total
        nb.arrays = {}
        nb._pins = []
        for k, a in batch.arrays.items():
            p = C.c_void_p()
            capi.check(self.lib.brb_host_alloc(max(1, a.nbytes), C.byref(p)), "brb_host_alloc")
            nb._pins.append(p)
            buf = (C.c_uint8 * max(1, a.nbytes)).from_address(p.value)
            na = np.frombuffer(buf, dtype=a.dtype, count=a.size).reshape(a.shape)
            na[...] = a
            nb.arrays[k] = na
            if hasattr(nb.desc, k):
                setattr(nb.desc, k, capi.ptr(na))
        return nb

    def pinned_free(self, nb):
        for p in nb._pins:
            self.lib.brb_host_free(p)
        nb._pins = []

    def stats(self):
        s = capi.Stats()
        capi.check(self.lib.brb_ctx_stats(self._ctx, C.byref(s)), "brb_ctx_stats")
        return s

    def lookup(self, params, refid, strand, qs, qe, status, cap):
        n = len(refid)
        hits = np.zeros((n, cap), dtype=np.uint32)
        nh = np.zeros(n, dtype=np.uint32)
        capi.check(self.lib.brb_lookup(self._ctx, C.byref(params), n, capi.ptr(refid), capi.ptr(strand), capi.ptr(qs), capi.ptr(qe),
                                       capi.ptr(status), cap, capi.ptr(hits.reshape(-1)), capi.ptr(nh)), "brb_lookup")
        return hits, nh

    def merge_cigars(self, real_off, real, ideal_off, ideal):
        n = len(real_off) - 1
        caps = (np.diff(real_off.astype(np.int64)) + np.diff(ideal_off.astype(np.int64)) + 2)
        out_off = np.concatenate([[0], np.cumsum(caps)]).astype(np.uint32)
        out_len = np.zeros(n, dtype=np.uint32)
        out = np.zeros(max(1, int(out_off[-1])), dtype=np.uint32)
        real = real if len(real) else np.zeros(1, np.uint32)
        ideal = ideal if len(ideal) else np.zeros(1, np.uint32)
        capi.check(self.lib.brb_merge_cigars(self._ctx, n, capi.ptr(real_off), capi.ptr(real), capi.ptr(ideal_off), capi.ptr(ideal),
                                             capi.ptr(out_off), capi.ptr(out_len), capi.ptr(out)), "brb_merge_cigars")
        return [out[int(o):int(o) + int(l)].copy() for o, l in zip(out_off[:-1], out_len)]

    def bgzf_compress(self, data):
        """bytes -> BGZF blocks (bytes), compressed on the device."""
        out = C.c_void_p(); n = C.c_uint64()
        capi.check(self.lib.brb_bgzf_compress(self._ctx, bytes(data), len(data), C.byref(out), C.byref(n)), "brb_bgzf_compress")
        return C.string_at(out.value, n.value) if n.value else b""

    def bgzf_inflate(self, data):
        """complete BGZF blocks (bytes) -> their inflated bytes, decoded and CRC-checked on the device."""
        out = C.c_void_p(); n = C.c_uint64()
        capi.check(self.lib.brb_bgzf_inflate(self._ctx, bytes(data), len(data), C.byref(out), C.byref(n)), "brb_bgzf_inflate")
        return C.string_at(out.value, n.value) if n.value else b""

    def emit_bgzf(self, params, raw, rec_off, rec_len, units, ctx=None):
        """Output records of the resident result (last project() on this engine's ctx, or `ctx`), assembled and BGZF-compressed on the
        device: raw = the bundle's input record bodies (bytes), rec_off / rec_len per alignment, units = (n, 4) uint32 in output
        order (first row, HI, NH, MAPQ | primary << 8).  Returns (BGZF bytes, uncompressed size)."""
        raw = np.frombuffer(bytes(raw) + b"\0", dtype=np.uint8)
        rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64); rec_len = np.ascontiguousarray(rec_len, dtype=np.uint32)
        units = np.ascontiguousarray(units, dtype=np.uint32).reshape(-1)
        ei = capi.EmitIn(raw.ctypes.data, len(raw) - 1, capi.ptr(rec_off if len(rec_off) else np.zeros(1, np.uint64)),
                         capi.ptr(rec_len if len(rec_len) else np.zeros(1, np.uint32)), len(units) // 4, capi.ptr(units if len(units) else np.zeros(4, np.uint32)))
        out = C.c_void_p(); n = C.c_uint64(); rawn = C.c_uint64()
        capi.check(self.lib.brb_emit_bgzf(ctx._ctx if ctx is not None else self._ctx, C.byref(params), C.byref(ei), C.byref(out), C.byref(n), C.byref(rawn)), "brb_emit_bgzf")
        return (C.string_at(out.value, n.value) if n.value else b""), rawn.value

    def move_result(self, to_ctx):
        capi.check(self.lib.brb_ctx_move_result(self._ctx, to_ctx._ctx), "brb_ctx_move_result")

    def ksw_extend(self, queries, targets, cigar_cap=256):
        n = len(queries)
        q_off = np.concatenate([[0], np.cumsum([len(q) for q in queries])]).astype(np.uint64)
        t_off = np.concatenate([[0], np.cumsum([len(t) for t in targets])]).astype(np.uint64)
        q = np.ascontiguousarray(np.concatenate(queries).astype(np.uint8)) if n else np.zeros(1, np.uint8)
        t = np.ascontiguousarray(np.concatenate(targets).astype(np.uint8)) if n else np.zeros(1, np.uint8)
        mx = np.zeros(n, np.int32); sc = np.zeros(n, np.int32); mq = np.zeros(n, np.int32); mt = np.zeros(n, np.int32)
        zd = np.zeros(n, np.uint8); nc = np.zeros(n, np.uint32); cg = np.zeros((n, cigar_cap), np.uint32)
        capi.check(self.lib.brb_ksw_extend(self._ctx, n, capi.ptr(q_off), capi.ptr(q), capi.ptr(t_off), capi.ptr(t), capi.ptr(mx),
                                           capi.ptr(sc), capi.ptr(mq), capi.ptr(mt), capi.ptr(zd), cigar_cap, capi.ptr(nc),
                                           capi.ptr(cg.reshape(-1))), "brb_ksw_extend")
        return dict(max=mx, score=sc, max_q=mq, max_t=mt, zdropped=zd, cigars=[cg[i, :min(int(nc[i]), cigar_cap)].copy() for i in range(n)], n_cigar=nc)
