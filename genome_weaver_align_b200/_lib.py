"""ctypes binding of libgwalign.so (include/gw_align.h). The library is CUDA only: there is no CPU path behind these calls,
and importing fails loudly when the shared object has not been built (run __graft_entry__.build() or `make -C csrc`)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "libgwalign.so")

GW_OK, GW_E_ARG, GW_E_IO, GW_E_CUDA, GW_E_CAPACITY = 0, -1, -2, -3, -4
READ_MAPPED, READ_UNMAPPED, READ_REF_ABORT, READ_UNSUPPORTED = 0, 1, 2, 3
CHR_NULL, CHR_EMPTY, CHR_STAR = -1, -2, -3


class Config(C.Structure):
    """gw_config_t == AlignmentConfig + AlignmentScoreConfig of the reference"""
    _fields_ = [("k", C.c_float), ("num_split_allowed", C.c_int32), ("match_score", C.c_int32), ("mismatch_penalty", C.c_int32),
                ("split_open_penalty", C.c_int32), ("indel_end_skip", C.c_int32), ("band_width", C.c_int32), ("report_type", C.c_int32),
                ("top_l", C.c_int32), ("strategy", C.c_int32), ("extended", C.c_int32)]


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("fm_searches", "sw_calls", "cutoffs", "filtered", "occ_lookups_scan", "occ_lookups_search",
                                          "occ_lookups_locate", "rounds", "gpu_launches")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


HIT_DTYPE = np.dtype([("read_index", "<u4"), ("chr_id", "<i4"), ("pos", "<i4"), ("match_length", "<i4"), ("diff", "<i4"), ("num_hits", "<i4"),
                      ("q_start", "<i4"), ("q_end", "<i4"), ("cigar_off", "<u4"), ("cigar_len", "<u4"), ("strand", "i1"), ("is_head", "u1"),
                      ("status", "u1"), ("reserved", "u1")])
assert HIT_DTYPE.itemsize == 44

# every symbol include/gw_align.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "gw_last_error": (C.c_char_p, []),
    "gw_index_build": (C.c_int, [C.c_char_p, C.c_int]),
    "gw_index_open": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(_P)]),
    "gw_index_locate_table": (C.c_int, [_P, C.c_int]),
    "gw_index_locate_step": (C.c_int, [_P]),
    "gw_index_from_sequence": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(_P)]),
    "gw_index_close": (None, [_P]),
    "gw_index_text_size": (C.c_int64, [_P]),
    "gw_index_num_contigs": (C.c_int32, [_P]),
    "gw_index_contig_name": (C.c_char_p, [_P, C.c_int32]),
    "gw_index_contig_length": (C.c_int64, [_P, C.c_int32]),
    "gw_index_contig_offset": (C.c_int64, [_P, C.c_int32]),
    "gw_index_device": (C.c_int, [_P]),
    "gw_config_default": (None, [C.POINTER(Config)]),
    "gw_aligner_create": (C.c_int, [_P, C.POINTER(Config), C.POINTER(_P)]),
    "gw_aligner_destroy": (None, [_P]),
    "gw_aligner_set_option": (C.c_int, [_P, C.c_char_p, C.c_int64]),
    "gw_align_batch": (C.c_int, [_P, _P, _P, C.c_uint32, _P, C.POINTER(C.c_uint64), _P, C.POINTER(C.c_uint64), C.POINTER(Counters)]),
    "gw_align_batch_fetch": (C.c_int, [_P, _P, C.POINTER(C.c_uint64), _P, C.POINTER(C.c_uint64)]),
    "gw_align_batch_device": (C.c_int, [_P, _P, _P, C.c_uint32, C.c_uint64, C.c_uint32, _P, C.c_uint64, _P, C.c_uint64, C.POINTER(C.c_uint64),
                                        C.POINTER(C.c_uint64), C.POINTER(Counters)]),
    "gw_aligner_read_counters": (C.c_int, [_P, C.c_uint32, C.c_uint32, _P]),
    "gw_aligner_last_timing": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "gw_sam_format": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, C.c_uint32, _P, C.c_uint64, _P, _P, C.POINTER(C.c_uint64), _P]),
    "gw_sam_header": (C.c_int, [_P, _P, C.POINTER(C.c_uint64)]),
    "gw_gather_benchmark": (C.c_int, [_P, C.c_int, C.c_uint64, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "gw_rank_acgtn": (C.c_int, [_P, C.c_int, _P, C.c_uint32, _P]),
    "gw_locate": (C.c_int, [_P, C.c_int, _P, C.c_uint32, _P]),
    "gw_quick_scan": (C.c_int, [_P, _P, _P, C.c_uint32, C.c_int, C.c_int, _P]),
    "gw_verify_pairs": (C.c_int, [C.c_int, _P, _P, _P, _P, C.c_uint32, C.c_int, _P, _P, C.c_uint32]),
    "gw_verify_pairs_noalloc": (C.c_int, [C.c_int, _P, _P, _P, _P, _P, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, _P, _P, C.c_uint32]),
    "gw_staircase_table": (C.c_int, [C.c_int, C.c_int, C.c_int, _P, C.c_int]),
    "gw_cyclic_suffix_array": (C.c_int, [_P, C.c_uint64, C.c_int, _P]),
    "gw_max_edit_distance_c": (C.c_int, [C.c_float, C.c_int]),
}

_lib = None


class GwError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libgwalign error %d: %s" % (code, msg))
        self.code = code


def lib():
    """loads libgwalign.so; raises if it is missing (the package has no other implementation to fall back to)"""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise ImportError("libgwalign.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(or make -C genome_weaver_align_b200/csrc); genome_weaver_align_b200 has no non-CUDA implementation")
        L = C.CDLL(SO_PATH)
        for name, (res, args) in SYMBOLS.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != GW_OK:
        raise GwError(rc, lib().gw_last_error().decode(errors="replace"))


def ptr(a):
    """numpy array / bytes -> void*"""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return C.cast(C.c_char_p(a), C.c_void_p).value


def concat(strings):
    """list of str/bytes -> (bytes, uint64 offsets[n+1])"""
    bs = [s.encode() if isinstance(s, str) else bytes(s) for s in strings]
    off = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        off[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64)
    return b"".join(bs), off
