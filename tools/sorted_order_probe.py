#!/usr/bin/env python
"""on the GPU box: how much would the short-read kernels gain if alignments came in genomic order?  The C2 bundle with its read-name
groups permuted by (refid, start of the first alignment) -- still a valid bundle (groups intact) -- against the original order, and with
only BATCHES of 32 groups' worth permuted (what a device-side batch sort could do)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bramble_b200 import workloads, engine, pack, capi
W = workloads.build('c2', 1000, int(sys.argv[1]) if len(sys.argv) > 1 else 1000000)
B = W.B
A = B.arrays
G = int(B.desc.n_groups); n = int(B.desc.n_aln)
go = A["group_off"].astype(np.int64)
first = go[:-1]
key = A["refid"].astype(np.int64)[first] * (1 << 32) + A["exon_st"].astype(np.int64)[A["exon_off"].astype(np.int64)[first]]

def permuted(order):
    """bundle with its groups in `order`"""
    parts = [B.slice_groups(int(g), int(g) + 1) for g in order] if len(order) < 2000 else None
    # vectorised: alignment permutation
    cnt = np.diff(go)[order]
    starts = go[:-1][order]
    idx = np.repeat(starts, cnt) + (np.arange(int(cnt.sum())) - np.repeat(np.cumsum(cnt) - cnt, cnt))
    inv = np.empty(n, np.int64); inv[idx] = np.arange(n)
    nb = pack.Batch(); X = nb.arrays
    X["group_off"] = np.concatenate([[0], np.cumsum(cnt)]).astype(np.uint32)
    X["refid"] = A["refid"][idx].copy(); X["strand"] = A["strand"][idx].copy()
    m = A["mate_idx"].astype(np.int64)[idx]; X["mate_idx"] = np.where(m >= 0, inv[np.maximum(m, 0)], -1).astype(np.int32)
    for off, cols in (("exon_off", ("exon_st", "exon_en")), ("cigar_off", ("cigar",))):
        o = A[off].astype(np.int64); c = (o[1:] - o[:-1])[idx]
        src = np.repeat(o[:-1][idx], c) + (np.arange(int(c.sum())) - np.repeat(np.cumsum(c) - c, c))
        X[off] = np.concatenate([[0], np.cumsum(c)]).astype(np.uint32)
        for k in cols: X[k] = A[k][src].copy()
    nb.desc.n_aln, nb.desc.n_groups = n, G
    for k, a in X.items(): setattr(nb.desc, k, capi.ptr(a))
    return nb

orders = {"original": np.arange(G), "groups sorted by position": np.argsort(key, kind="stable")}
blk = 16                                                     # ~32 alignments
nb_ = (G + blk - 1) // blk
bkey = key[::blk][:nb_]
orders["blocks of 16 groups sorted by position"] = (np.argsort(bkey, kind="stable")[:, None] * blk + np.arange(blk)[None, :]).reshape(-1)
orders["blocks of 16 groups sorted by position"] = orders["blocks of 16 groups sorted by position"][orders["blocks of 16 groups sorted by position"] < G]
with engine.Engine(W.ix) as eng:
    for name, order in orders.items():
        b = permuted(order)
        eng.upload(b)
        for _ in range(3): eng.run_resident(W.params)
        eng.timer_start()
        for _ in range(10): eng.run_resident(W.params)
        ms = eng.timer_stop() / 10
        st = eng.stats()
        print(f"{name:42s} {ms:.4f} ms/step  front {st.ms_lookup:.4f}  back {st.ms_pair:.4f}  rows {st.n_rec}")
