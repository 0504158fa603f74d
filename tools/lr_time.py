import sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import common
from bramble_b200 import engine, capi
for mode, n in (("lr", 6000), ("lr_nofasta", 6000), ("lrhq", 3000)):
    wl = common.make_workload(mode, 5, n_names=n, n_genes=300, n_chrom=3)
    ix, p = wl["index"], wl["params"]
    with engine.Engine(ix) as eng:
        for b in wl["batches"]:
            eng.upload(b)
            for _ in range(3): eng.run_resident(p)
            s = eng.stats()
            print(mode, "aln", b.desc.n_aln, "cand", s.n_candidates, "matches", s.n_matches, "rescue", s.n_rescue, "rec", s.n_rec,
                  "ms", round(s.kernel_ms, 3), "lookup", round(s.ms_lookup,3), "chain", round(s.ms_chain,3), "rescue", round(s.ms_rescue,3),
                  "build", round(s.ms_finalize,3), "back", round(s.ms_pair,3), "launches", s.n_launches, "Maln/s", round(b.desc.n_aln / s.kernel_ms / 1e3, 2))
