"""ctypes binding of tests/hostsim/libgwhostsim.so — TEST INFRASTRUCTURE ONLY.
The host simulator runs the per-thread kernel bodies of the product (compiled with g++) sequentially, so the automaton logic can be
compared with the oracle on a machine without a GPU. The product itself never loads it."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_DIR = os.path.join(ROOT, "tests", "hostsim")
_lib = None

DEFAULT_CFG = dict(k=0.1, num_split_allowed=1, match_score=1, mismatch_penalty=3, split_open_penalty=11,
                   indel_end_skip=5, band_width=31, report_type=0, top_l=5, extended=0)


def lib():
    global _lib
    if _lib is None:
        subprocess.check_call(["make", "-s", "-C", _DIR])
        L = C.CDLL(os.path.join(_DIR, "libgwhostsim.so"))
        L.gwsim_last_error.restype = C.c_char_p
        L.gwsim_open.restype = C.c_void_p
        L.gwsim_open.argtypes = [C.c_char_p]
        L.gwsim_close.argtypes = [C.c_void_p]
        L.gwsim_rank_acgtn.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.POINTER(C.c_uint32)]
        L.gwsim_locate.restype = C.c_uint32
        L.gwsim_locate.argtypes = [C.c_void_p, C.c_int, C.c_uint32]
        L.gwsim_align.restype = C.c_void_p
        L.gwsim_align.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_uint32, C.c_char_p, C.c_void_p, C.c_char_p, C.c_void_p,
                                  C.c_char_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                  C.POINTER(C.c_uint32)]
        L.gwsim_free.argtypes = [C.c_void_p]
        L.gwsim_verify_pair.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_uint32),
                                        C.POINTER(C.c_int)]
        _lib = L
    return _lib


def concat(strings):
    """list of str -> (bytes, uint64 offsets array of len n+1)"""
    off = np.zeros(len(strings) + 1, dtype=np.uint64)
    if strings:
        off[1:] = np.cumsum([len(s) for s in strings])
    return "".join(strings).encode(), off


class SimIndex:
    def __init__(self, fasta_path):
        self.h = lib().gwsim_open(fasta_path.encode())
        if not self.h:
            raise RuntimeError("hostsim: " + lib().gwsim_last_error().decode())

    def rank_acgtn(self, which, pos):
        out = (C.c_uint32 * 5)()
        lib().gwsim_rank_acgtn(self.h, which, pos, out)
        return list(out)

    def locate(self, which, row):
        return lib().gwsim_locate(self.h, which, row)

    def align(self, names, seqs, quals=None, slab_words=4096, **cfg):
        c = dict(DEFAULT_CFG)
        c.update(cfg)
        arr = (C.c_double * 10)(c["k"], c["num_split_allowed"], c["match_score"], c["mismatch_penalty"], c["split_open_penalty"],
                                c["indel_end_skip"], c["band_width"], c["report_type"], c["top_l"], c["extended"])
        n = len(seqs)
        b, boff = concat(seqs)
        nm, noff = concat(names)
        q, qoff = (None, None) if quals is None else concat(quals)
        status = (C.c_int * max(n, 1))()
        counters = (C.c_int * max(4 * n, 1))()
        lines = (C.c_int * max(n, 1))()
        tiers = C.c_uint32(0)
        p = lib().gwsim_align(self.h, arr, n, b, boff.ctypes.data, nm, noff.ctypes.data, q, None if qoff is None else qoff.ctypes.data,
                              slab_words, status, counters, lines, C.byref(tiers))
        if not p:
            raise RuntimeError("hostsim: " + lib().gwsim_last_error().decode())
        text = C.string_at(p).decode()
        lib().gwsim_free(p)
        all_lines = text.split("\n")[:-1] if text else []
        per_read, k = [], 0
        for i in range(n):
            per_read.append(all_lines[k:k + lines[i]])
            k += lines[i]
        return dict(sam=per_read, status=list(status)[:n], counters=[list(counters[4 * i:4 * i + 4]) for i in range(n)],
                    overflow_reads=tiers.value)

    def close(self):
        if self.h:
            lib().gwsim_close(self.h)
            self.h = None


_OPS = "MIDNSHPX"


def verify_pair(ref, query, band_k):
    cig = (C.c_uint32 * 64)()
    pos, nm, n = C.c_int(), C.c_int(), C.c_int()
    ok = lib().gwsim_verify_pair(ref.replace(" ", "").encode(), query.replace(" ", "").encode(), band_k, C.byref(pos), C.byref(nm), cig, C.byref(n))
    if not ok:
        return None
    return dict(cigar="".join("%d%s" % (cig[i] >> 4, _OPS[cig[i] & 7]) for i in range(n.value)), pos=pos.value, nm=nm.value)
