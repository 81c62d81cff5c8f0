"""ctypes binding of the CPU oracle (oracle/libgworacle.so). TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs, never by the product package."""
import ctypes as C
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(ROOT, "oracle", "libgworacle.so")
_lib = None

DEFAULT_CFG = dict(k=0.1, num_split_allowed=1, match_score=1, mismatch_penalty=3, split_open_penalty=11,
                   indel_end_skip=5, band_width=31, report_type=0, top_l=5, extended=0)


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        L.gwo_last_error.restype = C.c_char_p
        L.gwo_index_from_sequence.restype = C.c_void_p
        L.gwo_index_from_sequence.argtypes = [C.c_char_p, C.c_char_p]
        L.gwo_index_load.restype = C.c_void_p
        L.gwo_index_load.argtypes = [C.c_char_p]
        L.gwo_index_build.argtypes = [C.c_char_p]
        L.gwo_index_free.argtypes = [C.c_void_p]
        L.gwo_free.argtypes = [C.c_void_p]
        L.gwo_index_text_size.restype = C.c_longlong
        L.gwo_index_text_size.argtypes = [C.c_void_p]
        L.gwo_sam_header.restype = C.c_void_p
        L.gwo_sam_header.argtypes = [C.c_void_p]
        L.gwo_align.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.c_int, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int,
                                C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                C.POINTER(C.c_ulonglong), C.POINTER(C.c_double)]
        L.gwo_align_block_detailed.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_int),
                                               C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.gwo_backward_search.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_longlong, C.c_longlong, C.POINTER(C.c_longlong)]
        L.gwo_rank_acgtn.argtypes = [C.c_void_p, C.c_int, C.c_longlong, C.POINTER(C.c_longlong)]
        L.gwo_locate.restype = C.c_longlong
        L.gwo_locate.argtypes = [C.c_void_p, C.c_int, C.c_longlong]
        L.gwo_bwt_string.restype = C.c_void_p
        L.gwo_bwt_string.argtypes = [C.c_void_p, C.c_int]
        L.gwo_cyclic_sais_int.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_int)]
        L.gwo_fast_count_acgtn.argtypes = [C.c_char_p, C.c_longlong, C.c_longlong, C.POINTER(C.c_longlong)]
        L.gwo_fast_count.restype = C.c_longlong
        L.gwo_fast_count.argtypes = [C.c_char_p, C.c_int, C.c_longlong, C.c_longlong]
        L.gwo_query_mask64.restype = C.c_ulonglong
        L.gwo_query_mask64.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.gwo_staircase_table.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_ulonglong), C.c_int]
        L.gwo_compress_roundtrip.argtypes = [C.c_char_p, C.c_int]
        L.gwo_bitvector_lshift.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_char_p, C.c_int]
        L.gwo_cigar_merge.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int]
        L.gwo_cigar_build.argtypes = [C.POINTER(C.c_int), C.c_char_p, C.c_int, C.c_char_p, C.c_int]
        L.gwo_ssa_roundtrip.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.POINTER(C.c_int)]
        L.gwo_load_acgt_string.restype = C.c_void_p
        L.gwo_load_acgt_string.argtypes = [C.c_char_p]
        _lib = L
    return _lib


def _take_str(ptr):
    s = C.string_at(ptr).decode()
    lib().gwo_free(ptr)
    return s


class OracleIndex:
    def __init__(self, handle):
        if not handle:
            raise RuntimeError("oracle: " + lib().gwo_last_error().decode())
        self.h = handle

    @classmethod
    def from_sequence(cls, name, seq):
        return cls(lib().gwo_index_from_sequence(name.encode(), seq.encode()))

    @classmethod
    def load(cls, fasta_path):
        return cls(lib().gwo_index_load(fasta_path.encode()))

    @staticmethod
    def build_files(fasta_path):
        if lib().gwo_index_build(fasta_path.encode()) != 0:
            raise RuntimeError("oracle: " + lib().gwo_last_error().decode())

    def text_size(self):
        return lib().gwo_index_text_size(self.h)

    def sam_header(self):
        return _take_str(lib().gwo_sam_header(self.h))

    def bwt(self, which):
        return _take_str(lib().gwo_bwt_string(self.h, which))

    def backward_search(self, which, ch, lb, ub):
        out = (C.c_longlong * 2)()
        lib().gwo_backward_search(self.h, which, ch, lb, ub, out)
        return out[0], out[1]

    def rank_acgtn(self, which, i):
        out = (C.c_longlong * 5)()
        lib().gwo_rank_acgtn(self.h, which, i, out)
        return list(out)

    def locate(self, which, sa_index):
        return lib().gwo_locate(self.h, which, sa_index)

    def align(self, names, seqs, quals=None, nthreads=1, **cfg):
        """returns dict(sam=[per-read list of SAM lines], status, counters (n x 4), occ_lookups, elapsed)"""
        c = dict(DEFAULT_CFG)
        c.update(cfg)
        arr = (C.c_double * 10)(c["k"], c["num_split_allowed"], c["match_score"], c["mismatch_penalty"], c["split_open_penalty"],
                                c["indel_end_skip"], c["band_width"], c["report_type"], c["top_l"], c["extended"])
        n = len(seqs)
        status = (C.c_int * max(n, 1))()
        counters = (C.c_int * max(4 * n, 1))()
        lines = (C.c_int * max(n, 1))()
        occ = C.c_ulonglong(0)
        el = C.c_double(0)
        out = C.c_void_p()
        rc = lib().gwo_align(self.h, arr, n, "\n".join(names).encode(), "\n".join(seqs).encode(),
                             None if quals is None else "\n".join(quals).encode(), nthreads, C.byref(out), status, counters, lines,
                             C.byref(occ), C.byref(el))
        if rc != 0:
            raise RuntimeError("oracle: " + lib().gwo_last_error().decode())
        text = _take_str(out.value)
        all_lines = text.split("\n")[:-1] if text else []
        per_read, p = [], 0
        for i in range(n):
            per_read.append(all_lines[p:p + lines[i]])
            p += lines[i]
        return dict(sam=per_read, status=list(status)[:n], counters=[list(counters[4 * i:4 * i + 4]) for i in range(n)],
                    occ_lookups=occ.value, elapsed=el.value)

    def close(self):
        if self.h:
            lib().gwo_index_free(self.h)
            self.h = None


def align_block_detailed(ref, query, k):
    buf = C.create_string_buffer(4096)
    pos, nm, score = C.c_int(), C.c_int(), C.c_int()
    ok = lib().gwo_align_block_detailed(ref.replace(" ", "").encode(), query.replace(" ", "").encode(), k, buf, 4096,
                                         C.byref(pos), C.byref(nm), C.byref(score))
    if not ok:
        return None
    return dict(cigar=buf.value.decode(), pos=pos.value, nm=nm.value, score=score.value)


def parse_sam(line):
    f = line.split("\t")
    tags = {t.split(":")[0]: t.split(":", 2)[2] for t in f[11:]}
    return dict(name=f[0], flag=int(f[1]), chr=f[2], start=int(f[3]), mapq=int(f[4]), cigar=f[5], mchr=f[6], mstart=int(f[7]),
                isize=int(f[8]), seq=f[9], qual=f[10], nm=int(tags["NM"]) if "NM" in tags else None, tags=tags,
                strand="-" if int(f[1]) & 0x10 else "+")


def revcomp(s):
    return s[::-1].translate(str.maketrans("ACGTNacgtn", "TGCANtgcan"))
