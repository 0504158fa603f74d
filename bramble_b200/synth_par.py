"""Bench-scale long-read bundles.  synth.make_long_reads builds its records one by one in Python (about 1.5 ms per
ONT-like read), so a bundle of the reference's size (>= 100 000 alignments, src/bramble.cpp:362) is generated in chunks
on worker processes (fork: the annotation and the genome are shared, not pickled), each chunk packed into the engine's
structure-of-arrays by pack.pack_batches, and the chunks concatenated.  Seeded per chunk: the result does not depend
on the number of workers."""
import multiprocessing as mp
import os

import numpy as np

from . import pack, synth

_G = {}


def _chunk(args):
    k, seed, n, params, median_len, want_recs = args
    ann, genome, ix = _G["ann"], _G["genome"], _G["ix"]
    recs = synth.make_long_reads(ann, genome, seed, n, params=params, median_len=median_len, name_prefix=f"L{k:04d}_")
    bs = pack.pack_batches(recs, ix, ann.chroms, long_reads=True, use_fasta=genome is not None, chunk_size=1 << 30)
    return k, (bs[0].arrays if bs else None), (recs if want_recs else None)


def make_long_batch(ann, genome, ix, seed, n_names, params=None, median_len=1200, chunk=2000, procs=None, keep_records=1):
    """-> (Batch of all chunks, record dicts of the first `keep_records` chunks in BAM order -- the CPU baseline's sample)."""
    jobs = []
    k = 0
    left = n_names
    while left > 0:
        n = min(chunk, left)
        jobs.append((k, seed + 1000 * k, n, params, median_len, k < keep_records))
        left -= n
        k += 1
    _G.update(ann=ann, genome=genome, ix=ix)
    procs = procs or max(1, min(len(jobs), len(os.sched_getaffinity(0))))
    if procs > 1:
        with mp.get_context("fork").Pool(procs) as pool:
            out = pool.map(_chunk, jobs, chunksize=1)
    else:
        out = [_chunk(j) for j in jobs]
    _G.clear()
    out.sort(key=lambda x: x[0])
    recs = [r for _, _, rr in out if rr for r in rr]
    return pack.concat_batches([a for _, a, _ in out if a is not None]), recs
