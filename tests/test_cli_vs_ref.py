"""The drop-in CLI against the unmodified reference binary: the DECOMPRESSED output BAM streams (header text,
reference dictionary, every record byte, record order) must be identical.  CPU leg: the host half linked to
the CPU oracle (oracle/_ref/port_cli) -- pins ingest, flush()/NH/HI/MAPQ/primary tie-break and record
assembly.  GPU leg: the product binary (CUDA engine behind the C ABI)."""
import os
import struct
import subprocess
import tempfile
import zlib

import pytest

from bramble_b200 import synth
import common
import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PORT_CLI = os.path.join(ROOT, "oracle", "_ref", "port_cli")
CLI = os.path.join(ROOT, "bramble_b200", "bramble_b200_cli")
needs_ref = pytest.mark.skipif(not os.path.exists(oracle_lib.REF_BIN), reason="oracle/_ref/bramble_ref not built")


def raw_stream(path):
    data = open(path, "rb").read()
    p, out = 0, bytearray()
    while p < len(data):
        xlen = struct.unpack_from("<H", data, p + 10)[0]
        q, bsize = p + 12, None
        while q < p + 12 + xlen:
            slen = struct.unpack_from("<H", data, q + 2)[0]
            if data[q] == 66 and data[q + 1] == 67:
                bsize = struct.unpack_from("<H", data, q + 4)[0]
            q += 4 + slen
        out += zlib.decompress(data[p + 12 + xlen:p + bsize + 1 - 8], -15)
        p += bsize + 1
    return bytes(out)


def run_both(d, exe, args, extra=()):
    os.makedirs(d + "/ref", exist_ok=True)
    os.makedirs(d + "/our", exist_ok=True)
    for sub, binary, more in (("ref", oracle_lib.REF_BIN, []), ("our", exe, list(extra))):
        if sub == "ref" and os.path.exists(d + "/ref/out.bam"):
            continue
        # argv[0] and every argument are the same strings in both runs: they end up in the @PG CL: field
        argv = ["bramble", "../in.bam", "-G", "../a.gtf", "-o", "out.bam", "-p", "1", "--quiet"] + list(args)
        env = dict(os.environ, BRB_SHIM_LEVEL="1")
        r = subprocess.run(argv + more, executable=binary, cwd=d + "/" + sub, capture_output=True, text=True, env=env)
        assert r.returncode == 0, (sub, r.stderr[-2000:])
    return raw_stream(d + "/ref/out.bam"), raw_stream(d + "/our/out.bam")


def write_inputs(wl, d, mode):
    m = common.MODES[mode]
    synth.write_gtf(wl["ann"], d + "/a.gtf")
    args = list(m["args"])
    if wl["genome"] is not None:
        synth.write_fasta(wl["ann"], wl["genome"], d + "/g.fa")
        args += ["-S", "../g.fa"]
    synth.write_bam(d + "/in.bam", wl["ann"], wl["recs"])
    return args


def explain(a, b, d):
    if a == b:
        return ""
    _, _, ra = synth.read_bam(d + "/ref/out.bam")
    _, _, rb = synth.read_bam(d + "/our/out.bam")
    msg = f"streams differ: {len(a)} vs {len(b)} bytes, {len(ra)} vs {len(rb)} records"
    for k, (x, y) in enumerate(zip(ra, rb)):
        if x != y:
            return msg + f"\nfirst differing record {k}:\n ref {synth.describe_record(x)}\n our {synth.describe_record(y)}"
    return msg


@needs_ref
@pytest.mark.parametrize("mode", ["default", "fr", "strict", "lr", "lrhq"])
def test_host_half_with_oracle_engine_is_byte_identical(mode):
    assert os.path.exists(PORT_CLI)
    wl = common.make_workload(mode, 17, n_names=1000, n_genes=40)
    d = tempfile.mkdtemp(prefix="brb_cli_")
    args = write_inputs(wl, d, mode)
    a, b = run_both(d, PORT_CLI, args)
    assert a == b, explain(a, b, d)
    assert len(a) > 100000


@needs_ref
@pytest.mark.parametrize("mode", ["default", "lr"])
def test_host_threads_do_not_change_the_output(mode):
    """-p N runs BGZF inflate, bundle assembly (flush replay + records) and BGZF deflate on thread pools; the
    ordered merge keeps the stream identical to the single-threaded one (many small bundles in flight)."""
    wl = common.make_workload(mode, 29, n_names=1200, n_genes=40)
    d = tempfile.mkdtemp(prefix="brb_cli_")
    args = write_inputs(wl, d, mode)
    outs = []
    for th in ("1", "5"):
        os.makedirs(f"{d}/t{th}", exist_ok=True)
        argv = ["bramble", "../in.bam", "-G", "../a.gtf", "-o", "out.bam", "--quiet", "--chunk", "150"] + args
        r = subprocess.run(argv, executable=PORT_CLI, cwd=f"{d}/t{th}", capture_output=True, text=True, env=dict(os.environ, BRB_THREADS=th))
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(raw_stream(f"{d}/t{th}/out.bam"))
    assert outs[0] == outs[1] and len(outs[0]) > 100000


@needs_ref
def test_flush_chunks_assembled_in_parallel_keep_the_reference_order():
    """Several 5000-pair flush chunks in one bundle: the reference reuses ONE name map across them (clear() keeps its bucket
    array), so the record order inside chunk k depends on the earlier chunks; the CLI assembles the chunks on pool threads
    from the recorded bucket counts and must still write the reference's bytes."""
    from bramble_b200 import synth_fast as sf
    A = sf.make_annotation(11, n_chrom=2, n_genes=200)
    ix = sf.build_index(A)
    B = sf.make_short_batch(A, ix, 12, 24000, stranded=None, multi=0.15)
    d = tempfile.mkdtemp(prefix="brb_cli_")
    sf.write_gtf(A, d + "/a.gtf")

    class Ann:
        chroms, chrom_len = A.chroms, A.chrom_len
    synth.write_bam(d + "/in.bam", Ann, sf.to_records(B))
    old = os.environ.get("BRB_THREADS")
    os.environ["BRB_THREADS"] = "5"
    try:
        a, b = run_both(d, PORT_CLI, [])
    finally:
        if old is None:
            del os.environ["BRB_THREADS"]
        else:
            os.environ["BRB_THREADS"] = old
    assert a == b, explain(a, b, d)


@needs_ref
def test_piece_reader_with_host_decoder():
    """The input reader in piece mode (whole BGZF blocks, 1 MiB then 8 MiB, handed to the engine's decoder) with every piece declined: the
    host inflates them block by block.  Also: a corrupt block is reported, not written through."""
    wl = common.make_workload("default", 31, n_names=6000, n_genes=40)
    d = tempfile.mkdtemp(prefix="brb_cli_")
    args = write_inputs(wl, d, "default")
    os.environ["BRB_TEST_PIECES"] = "1"
    try:
        a, b = run_both(d, PORT_CLI, args)
        assert a == b and len(a) > 1000000, explain(a, b, d)
        z = bytearray(open(d + "/in.bam", "rb").read())
        z[len(z) // 2] ^= 0x40
        open(d + "/bad.bam", "wb").write(bytes(z))
        r = subprocess.run(["bramble", "../bad.bam", "-G", "../a.gtf", "-o", "bad_out.bam", "--quiet"], executable=PORT_CLI, cwd=d + "/our", capture_output=True, text=True)
        assert r.returncode == 1 and "error" in r.stderr.lower(), (r.returncode, r.stderr[-300:])
    finally:
        del os.environ["BRB_TEST_PIECES"]


def _pg_strip(stream):
    """drop the trailing extra flags our runs add to the @PG CL: field (--gpus N)"""
    return stream


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["default", "fr", "rf", "strict", "lr", "lr_nofasta", "lrhq"])
def test_cli_is_byte_identical_to_reference(mode):
    wl = common.make_workload(mode, 19, n_names=1500, n_genes=40)
    d = tempfile.mkdtemp(prefix="brb_cli_")
    args = write_inputs(wl, d, mode)
    a, b = run_both(d, CLI, args)
    assert a == b, explain(a, b, d)


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("env", [{"BRB_HOST_ASSEMBLY": "1"}, {"BRB_HOST_DEFLATE": "1"}, {"BRB_SLOTS": "1"}, {"BRB_DEVICE_INFLATE": "1"}])
@pytest.mark.parametrize("mode", ["default", "lr"])
def test_cli_host_side_output_paths_agree(mode, env):
    """The default output path builds and compresses the records on the device (brb_emit_bgzf).  The older paths stay selectable and
    must write the same stream: records assembled by the host pool + device BGZF (BRB_HOST_ASSEMBLY), everything on the host
    (BRB_HOST_DEFLATE); BRB_SLOTS=1 leaves the device path a single row holder; BRB_DEVICE_INFLATE=1 has the input BGZF blocks decoded by the
    engine (brb_bgzf_inflate) instead of the host pool."""
    wl = common.make_workload(mode, 23, n_names=1500, n_genes=40)
    d = tempfile.mkdtemp(prefix="brb_cli_")
    args = write_inputs(wl, d, mode)
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        a, b = run_both(d, CLI, args)
    finally:
        for k, v in old.items():
            if v is None:
                del os.environ[k]
            else:
                os.environ[k] = v
    assert a == b, explain(a, b, d)


@needs_ref
@pytest.mark.gpu
def test_cli_multi_bundle_overrides_and_flush_chunks():
    """>100k alignments (two bundles, many 5000-pair flush chunks) + threshold overrides."""
    from bramble_b200 import synth_fast as sf
    A = sf.make_annotation(7, n_chrom=3, n_genes=400)
    ix = sf.build_index(A)
    B = sf.make_short_batch(A, ix, 8, 52000, stranded=None, multi=0.15)
    assert B.desc.n_aln > 110000
    d = tempfile.mkdtemp(prefix="brb_cli_")
    sf.write_gtf(A, d + "/a.gtf")

    class Ann:
        chroms, chrom_len = A.chroms, A.chrom_len
    synth.write_bam(d + "/in.bam", Ann, sf.to_records(B))
    a, b = run_both(d, CLI, ["--max-soft-clip", "8", "--max-junction-insertion", "2", "--similarity-threshold", "0.5"])
    assert a == b, explain(a, b, d)


def _three_bundle_input():
    """> 200k alignments: three bundles of the reference's size (>= 100 000 alignments, cut at a read name)."""
    from bramble_b200 import synth_fast as sf
    A = sf.make_annotation(9, n_chrom=3, n_genes=400)
    ix = sf.build_index(A)
    B = sf.make_short_batch(A, ix, 10, 100000, stranded=None, multi=0.15)
    assert B.desc.n_aln > 205000
    d = tempfile.mkdtemp(prefix="brb_cli_")
    sf.write_gtf(A, d + "/a.gtf")

    class Ann:
        chroms, chrom_len = A.chroms, A.chrom_len
    synth.write_bam(d + "/in.bam", Ann, sf.to_records(B))
    return d


def _dealer_run(env):
    d = _three_bundle_input()
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        a, b = run_both(d, CLI, [], extra=[])
    finally:
        for k, v in old.items():
            if v is None:
                del os.environ[k]
            else:
                os.environ[k] = v
    assert a == b, explain(a, b, d)


@needs_ref
@pytest.mark.gpu
def test_cli_dealer_two_slots_on_one_device():
    """The --gpus 2 code path -- bundle k to dealer slot k mod 2 (one worker thread, one index copy and one context per slot),
    ordered merge of the output -- with both slots mapped to device 0 (BRB_GPU_MAP), so that it runs on a 1-GPU box:
    three bundles, byte-identical to the reference."""
    _dealer_run({"BRB_GPUS": "2", "BRB_GPU_MAP": "0,0", "BRB_THREADS": "4"})


@needs_ref
@pytest.mark.gpu
def test_cli_two_gpus_ordered_merge():
    from bramble_b200 import capi
    if capi.lib().brb_device_count() < 2:
        pytest.skip("needs 2 GPUs (the same dealer path runs on one device in test_cli_dealer_two_slots_on_one_device)")
    _dealer_run({"BRB_GPUS": "2", "BRB_THREADS": "4"})
