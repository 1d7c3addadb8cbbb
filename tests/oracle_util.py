"""ctypes wrapper of the oracle (oracle/_build/liboracle.so) -- the CPU
restatement of the reference used as the checker in tests, smoke() and the
cpu_baseline leg of bench.py.  Never used by the product."""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "_build", "liboracle.so")
SCHEMES_DIR = os.path.join(ROOT, "columba_b200", "schemes")

PART = {"uniform": 0, "static": 1, "dynamic": 2}
METRIC = {"hamming": 0, "edit": 1}


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "oracle"])
    return ORACLE_LIB


class Oracle:
    def __init__(self, base, sa_sparse=4, kmer=10, in_text=4, scheme=("named", "columba"), partition="dynamic",
                 metric="edit"):
        if not os.path.exists(ORACLE_LIB):
            build_oracle()
        self.lib = C.CDLL(ORACLE_LIB)
        self.lib.orc_open.restype = C.c_void_p
        self.lib.orc_open.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_char_p, C.c_char_p,
                                      C.c_int, C.c_int]
        self.lib.orc_close.argtypes = [C.c_void_p]
        self.lib.orc_map_read.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_uint64),
                                          C.POINTER(C.c_void_p)]
        self.lib.orc_free.argtypes = [C.c_void_p]
        self.lib.orc_occ4.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.POINTER(C.c_uint32)]
        self.lib.orc_extend.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        self.lib.orc_find_sa.argtypes = [C.c_void_p, C.c_uint32]
        self.lib.orc_find_sa.restype = C.c_uint32
        self.lib.orc_text_length.argtypes = [C.c_void_p]
        self.lib.orc_text_length.restype = C.c_uint32
        self.lib.orc_seq_name.argtypes = [C.c_void_p, C.c_uint32]
        self.lib.orc_seq_name.restype = C.c_char_p
        self.h = self.lib.orc_open(base.encode(), sa_sparse, kmer, in_text, scheme[0].encode(), scheme[1].encode(),
                                   SCHEMES_DIR.encode(), PART[partition], METRIC[metric])
        if not self.h:
            raise RuntimeError("oracle could not open " + base)
        self.counters = (C.c_uint64 * 17)()

    def close(self):
        if self.h:
            self.lib.orc_close(self.h)
            self.h = None

    def text_length(self):
        return self.lib.orc_text_length(self.h)

    def occ4(self, p, reverse=False):
        out = (C.c_uint32 * 4)()
        self.lib.orc_occ4(self.h, int(reverse), p, out)
        return list(out)

    def extend(self, parent, mode):
        a = (C.c_uint32 * 4)(*parent)
        out = (C.c_uint32 * 16)()
        self.lib.orc_extend(self.h, mode, a, out)
        return [list(out[i * 4:(i + 1) * 4]) for i in range(4)]

    def find_sa(self, row):
        return self.lib.orc_find_sa(self.h, row)

    def map_read(self, seq, k):
        """-> list of (strand, seq_id, pos1, dist, cigar)"""
        txt = C.c_void_p()
        n = self.lib.orc_map_read(self.h, seq.encode(), k, self.counters, C.byref(txt))
        if n < 0:
            raise RuntimeError("oracle failed")
        s = C.string_at(txt).decode()
        self.lib.orc_free(txt)
        out = []
        for line in s.splitlines():
            a = line.split("\t")
            out.append((int(a[0]), int(a[1]), int(a[2]), int(a[3]), a[4]))
        return out

    def tuples(self, names, seqs, k):
        out = []
        for nm, s in zip(names, seqs):
            for strand, sid, pos, dist, cg in self.map_read(s, k):
                out.append((nm, 16 if strand else 0, self.lib.orc_seq_name(self.h, sid).decode(), pos, dist, cg))
        out.sort()
        return out
