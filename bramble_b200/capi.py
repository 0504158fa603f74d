"""ctypes view of include/bramble_b200.h (the C ABI of the CUDA projection engine).

The structures here are shared with the test-side binding of the CPU oracle (tests/oracle_lib.py),
which exports the same data contract.  Nothing in this module computes anything: it loads
`libbramble_b200.so` (built in-tree by `__graft_entry__.build()`), and raises loudly when the
library or a CUDA device is missing -- there is no CPU implementation behind this ABI.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BRB_LIB") or os.path.join(_HERE, "libbramble_b200.so")     # BRB_LIB: tuning builds (tools/variants.py)

BRB_OK, BRB_E_NO_DEVICE, BRB_E_CUDA, BRB_E_ARG, BRB_E_NOMEM, BRB_E_UNSUPPORTED = 0, -1, -2, -3, -4, -5
REC_FIRST, REC_PAIRED, REC_SAMETX = 1, 2, 4
OPT_GENERIC_FRONT = 1
OPT_NO_FASTPATH = 2
OPT_ORDERED_ROWS = 3
OPT_PACKED_ROWS = 4
OPT_MERGE_WARP = 5
OPT_FLUSH_TAGS = 6

u8p, i8p = C.POINTER(C.c_uint8), C.POINTER(C.c_int8)
u32p, i32p = C.POINTER(C.c_uint32), C.POINTER(C.c_int32)
u64p, f64p = C.POINTER(C.c_uint64), C.POINTER(C.c_double)


class Params(C.Structure):
    _fields_ = [("max_clip", C.c_uint32), ("max_junc_ins", C.c_uint32), ("max_junc_gap", C.c_uint32),
                ("max_error_exon", C.c_uint32), ("similarity_threshold", C.c_float),
                ("long_reads", C.c_uint8), ("use_fasta", C.c_uint8), ("ignore_small_exons", C.c_uint8),
                ("filter_by_similarity", C.c_uint8)]


class IndexDesc(C.Structure):
    _fields_ = [("n_tables", C.c_uint32), ("table_off", u64p), ("st", i32p), ("en", i32p), ("tid", u32p),
                ("exon_id", u8p), ("pos_start", u32p), ("prev_st", u32p), ("prev_en", u32p),
                ("next_st", u32p), ("next_en", u32p), ("has_prev_next", u8p), ("transcript_len", u32p),
                ("n_chrom", C.c_uint32), ("genome_format", C.c_uint32), ("chrom_off", u64p), ("genome_codes", u8p)]


class BatchIn(C.Structure):
    _fields_ = [("n_aln", C.c_uint32), ("n_groups", C.c_uint32), ("group_off", u32p), ("refid", i32p),
                ("strand", i8p), ("exon_off", u32p), ("exon_st", u32p), ("exon_en", u32p),
                ("cigar_off", u32p), ("cigar", u32p), ("mate_idx", i32p), ("group_seq_off", u64p),
                ("seq_codes", u8p), ("group_name_off", u64p), ("name_bytes", u8p),
                ("pos0", i32p), ("cigar_cnt16", C.POINTER(C.c_uint16)), ("mate_delta16", C.POINTER(C.c_int16)), ("refid16", C.POINTER(C.c_int16)),
                ("n_cigar_total", C.c_uint64)]


class BatchOut(C.Structure):
    _fields_ = [("n_rec", C.c_uint32), ("n_cigar_words", C.c_uint64), ("group_rec_off", u32p),
                ("src_aln", u32p), ("tid", u32p), ("pos", u32p), ("tx_strand", i8p), ("flags", u8p),
                ("sim_score", f64p), ("clip_score", i32p), ("cigar_off", u32p), ("cigar", u32p),
                ("aln_rec_off", u32p), ("aln_rec_cnt", u32p), ("packed_rows", u32p),
                ("hi", i32p), ("nh", i32p), ("mapq", u8p), ("primary", u8p),
                ("batch_rec_off", u32p), ("aln_rec_cnt16", C.POINTER(C.c_uint16)), ("packed_row_words", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("n_launches", C.c_uint64), ("kernel_ms", C.c_float), ("ms_lookup", C.c_float),
                ("ms_chain", C.c_float), ("ms_rescue", C.c_float), ("ms_finalize", C.c_float),
                ("ms_pair", C.c_float), ("ms_merge", C.c_float), ("n_candidates", C.c_uint64),
                ("n_matches", C.c_uint64), ("n_rescue", C.c_uint64), ("n_rec", C.c_uint64),
                ("alg_bytes", C.c_uint64), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("alg_bytes_front", C.c_uint64), ("ksw_cells", C.c_uint64), ("host_syncs", C.c_uint32), ("reserved_", C.c_uint32), ("ksw_cells_run", C.c_uint64)]


_PTR = {np.dtype(np.uint8): u8p, np.dtype(np.int8): i8p, np.dtype(np.uint32): u32p, np.dtype(np.int32): i32p,
        np.dtype(np.uint64): u64p, np.dtype(np.float64): f64p, np.dtype(np.uint16): C.POINTER(C.c_uint16), np.dtype(np.int16): C.POINTER(C.c_int16)}


def ptr(a, dtype=None):
    """Pointer to a C-contiguous numpy array (None -> NULL).  The caller keeps `a` alive."""
    if a is None:
        return None
    if dtype is not None:
        assert a.dtype == np.dtype(dtype), (a.dtype, dtype)
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_PTR[a.dtype])


def resolve_params(long_reads=False, lr=False, lr_hq=False, strict=False, use_fasta=False, max_clip=None,
                   max_junc_ins=None, max_junc_gap=None, max_error_exon=None, similarity_threshold=None):
    """Threshold presets of the reference: ShortReadEvaluator::evaluate src/evaluate.cpp:1136-1175,
    LongReadEvaluator::evaluate :1177-1221; CLI overrides win (std::optional::value_or)."""
    long_reads = bool(long_reads or lr or lr_hq)
    if not long_reads:
        d = dict(clip=0 if strict else 5, ins=0, gap=0, sim=1.0, err=0)
    elif lr or not lr_hq:
        d = dict(clip=40, ins=40, gap=40, sim=0.60, err=35)
    else:
        d = dict(clip=5, ins=10, gap=10, sim=0.90, err=35)
    p = Params()
    p.max_clip = d["clip"] if max_clip is None else max_clip
    p.max_junc_ins = d["ins"] if max_junc_ins is None else max_junc_ins
    p.max_junc_gap = d["gap"] if max_junc_gap is None else max_junc_gap
    p.max_error_exon = d["err"] if max_error_exon is None else max_error_exon
    p.similarity_threshold = d["sim"] if similarity_threshold is None else similarity_threshold
    p.long_reads = int(long_reads)
    p.use_fasta = int(bool(use_fasta))
    p.ignore_small_exons = int(p.max_error_exon > 0)
    p.filter_by_similarity = int(np.float32(p.similarity_threshold) < np.float32(1.0))
    return p


class EmitIn(C.Structure):
    _fields_ = [("raw", C.c_void_p), ("raw_bytes", C.c_uint64), ("rec_off", u64p), ("rec_len", u32p), ("n_units", C.c_uint32), ("units", u32p)]


def _declare(lib):
    lib.brb_device_count.restype = C.c_int
    lib.brb_last_error.restype = C.c_char_p
    lib.brb_version.restype = C.c_char_p
    lib.brb_index_create.argtypes = [C.c_int, C.POINTER(IndexDesc), C.POINTER(C.c_void_p)]
    lib.brb_index_destroy.argtypes = [C.c_void_p]
    lib.brb_ctx_create.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
    lib.brb_ctx_destroy.argtypes = [C.c_void_p]
    lib.brb_project.argtypes = [C.c_void_p, C.POINTER(Params), C.POINTER(BatchIn), C.POINTER(BatchOut)]
    lib.brb_batch_upload.argtypes = [C.c_void_p, C.POINTER(BatchIn)]
    lib.brb_batch_read_exons.argtypes = [C.c_void_p, u32p, u32p, u32p, C.c_uint64, C.POINTER(C.c_uint64)]
    lib.brb_project_resident.argtypes = [C.c_void_p, C.POINTER(Params)]
    lib.brb_batch_download.argtypes = [C.c_void_p, C.POINTER(BatchOut)]
    lib.brb_ctx_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
    lib.brb_ctx_set_option.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.brb_ctx_timer_start.argtypes = [C.c_void_p]
    lib.brb_ctx_timer_stop.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
    lib.brb_lookup.argtypes = [C.c_void_p, C.POINTER(Params), C.c_uint32, i32p, i8p, u32p, u32p, u8p, C.c_uint32, u32p, u32p]
    lib.brb_merge_cigars.argtypes = [C.c_void_p, C.c_uint32, u32p, u32p, u32p, u32p, u32p, u32p, u32p]
    lib.brb_ksw_extend.argtypes = [C.c_void_p, C.c_uint32, u64p, u8p, u64p, u8p, i32p, i32p, i32p, i32p, u8p,
                                   C.c_uint32, u32p, u32p]
    lib.brb_bgzf_compress.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.brb_bgzf_inflate.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.brb_emit_bgzf.argtypes = [C.c_void_p, C.POINTER(Params), C.POINTER(EmitIn), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.brb_ctx_move_result.argtypes = [C.c_void_p, C.c_void_p]
    lib.brb_merge_cigars_static.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32] + [C.c_void_p] * 8     # device pointers
    lib.brb_sort_by_start.argtypes = [C.c_uint64, i32p, u32p]
    lib.brb_host_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
    lib.brb_host_free.argtypes = [C.c_void_p]
    for f in ("brb_index_create", "brb_index_destroy", "brb_ctx_create", "brb_ctx_destroy", "brb_project",
              "brb_batch_upload", "brb_project_resident", "brb_batch_download", "brb_ctx_stats", "brb_ctx_set_option", "brb_ctx_timer_start",
              "brb_ctx_timer_stop", "brb_lookup",
              "brb_merge_cigars", "brb_ksw_extend", "brb_sort_by_start", "brb_host_alloc", "brb_host_free", "brb_bgzf_compress", "brb_emit_bgzf", "brb_ctx_move_result", "brb_merge_cigars_static", "brb_bgzf_inflate", "brb_batch_read_exons"):
        getattr(lib, f).restype = C.c_int
    return lib


_lib = None


def lib():
    """The engine library.  Fails loudly if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(nvcc, sm_100a).  bramble_b200 has no CPU implementation.")
        _lib = _declare(C.CDLL(LIB_PATH))
    return _lib


class EngineError(RuntimeError):
    pass


def check(rc, what=""):
    if rc != BRB_OK:
        msg = lib().brb_last_error()
        raise EngineError(f"{what} failed with status {rc}: {msg.decode() if msg else ''}")
