#!/usr/bin/env python
"""on the GPU box: brb_bgzf_inflate on a C2 BAM sample (300k read names), whole file and 4 MB pieces, against zlib on one host core"""
import sys, os, time, zlib, struct
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bramble_b200 import workloads, engine, synth
import bench
W = workloads.build('c2', 1000, 300000)
rb = bench.RefBench(W, {"small": 2000, "cli": 300000})
z = open(rb.d + "/cli.bam", "rb").read()
print("file", len(z))
t = time.time(); parts = []; p = 0
while p < len(z):
    bs = struct.unpack_from("<H", z, p + 16)[0] + 1
    parts.append(zlib.decompress(z[p + 18:p + bs - 8], -15)); p += bs
t_host = time.time() - t
raw = b"".join(parts)
print("zlib 1 core: %.3f s, %d bytes" % (t_host, len(raw)))
with engine.Engine(W.ix) as eng:
    for rep in range(3):
        eng.timer_start(); t = time.time(); got = eng.bgzf_inflate(z); dt = time.time() - t; ms = eng.timer_stop()
        print("device whole file: %.4f s wall in python, %.2f ms on the stream (H2D + kernel + D2H; %.1f GB/s out) ok=%s" % (dt, ms, len(got) / ms / 1e6, got == raw))
    # pieces of ~4 MB at block boundaries
    cuts = [0]; p = 0
    while p < len(z):
        p += struct.unpack_from("<H", z, p + 16)[0] + 1
        if p - cuts[-1] >= (4 << 20) or p == len(z): cuts.append(p)
    for rep in range(2):
        t = time.time(); n = 0
        for a, b in zip(cuts[:-1], cuts[1:]): n += len(eng.bgzf_inflate(z[a:b]))
        dt = time.time() - t
        print("device %d pieces: %.4f s (%.1f GB/s out)" % (len(cuts) - 1, dt, n / dt / 1e9))
