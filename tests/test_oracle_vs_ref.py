"""CPU: pins the oracle.  The CPU restatement (oracle/libbrb_port.so) + the record assembly of tests/refcmp.py
must reproduce, record for record (tid, pos, flag, MAPQ, CIGAR, mate fields, NH, HI, AS), the output BAM of
the UNMODIFIED reference compiled into oracle/_ref/bramble_ref, in every mode."""
import os
import tempfile

import pytest

from bramble_b200 import synth
import common
import oracle_lib
import refcmp

needs_ref = pytest.mark.skipif(not os.path.exists(oracle_lib.REF_BIN), reason="oracle/_ref/bramble_ref not built (no reference sources)")


@needs_ref
@pytest.mark.parametrize("mode", ["default", "fr", "rf", "strict", "lr", "lr_nofasta", "lrhq"])
def test_port_matches_reference_binary(mode):
    m = common.MODES[mode]
    wl = common.make_workload(mode, 11, n_names=1200, n_genes=40, chunk_size=100000)
    d = tempfile.mkdtemp(prefix="brb_ref_")
    synth.write_gtf(wl["ann"], d + "/a.gtf")
    args = list(m["args"])
    if wl["genome"] is not None:
        synth.write_fasta(wl["ann"], wl["genome"], d + "/g.fa")
        args += ["-S", d + "/g.fa"]
    synth.write_bam(d + "/in.bam", wl["ann"], wl["recs"])
    oracle_lib.run_reference(d, d + "/in.bam", d + "/a.gtf", d + "/out.bam", args)
    text, refs, raw = synth.read_bam(d + "/out.bam")
    ix = wl["index"]
    assert [r[0] for r in refs] == ix.tid_names, "transcript (tid) order differs from the reference header"
    assert [r[1] for r in refs] == ix.tid_len
    rows = [(b, oracle_lib.port_project(ix, wl["params"], b)) for b in wl["batches"]]
    exp = refcmp.expected_records(rows, wl["recs"], m["long"])
    got = refcmp.parse_output_bam(raw, m["long"])
    assert len(raw) > 100
    refcmp.compare(exp, got)


@needs_ref
def test_port_matches_reference_micro_introns_and_shuffled_gtf():
    wl = common.make_workload("lr_nofasta", 13, n_names=800, n_genes=40, micro_introns=True)
    d = tempfile.mkdtemp(prefix="brb_ref_")
    synth.write_gtf(wl["ann"], d + "/a.gtf")
    synth.write_bam(d + "/in.bam", wl["ann"], wl["recs"])
    oracle_lib.run_reference(d, d + "/in.bam", d + "/a.gtf", d + "/out.bam", ["--lr"])
    _, refs, raw = synth.read_bam(d + "/out.bam")
    assert [r[0] for r in refs] == wl["index"].tid_names
    rows = [(b, oracle_lib.port_project(wl["index"], wl["params"], b)) for b in wl["batches"]]
    refcmp.compare(refcmp.expected_records(rows, wl["recs"], True), refcmp.parse_output_bam(raw, True))
