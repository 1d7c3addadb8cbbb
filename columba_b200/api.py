"""ctypes binding of the C ABI in include/columba_b200.h.

The product library is columba_b200/libcolumba_b200.so (nvcc, sm_100a).  There
is no CPU fallback: loading it works anywhere (so its exported symbols can be
checked), but every computing entry point needs a CUDA device and raises
ClbError otherwise.  `lib_path` exists so the CPU unit tests of the kernel
logic can point the same binding at tests/_work/libclb_hostemu.so.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_LIB = os.environ.get("COLUMBA_B200_LIB") or os.path.join(HERE, "libcolumba_b200.so")  # override: A/B runs of two builds
SCHEMES_DIR = os.path.join(HERE, "schemes")

N_COUNTERS, N_STATS, N_STAGES = 17, 11, 6  # CLB_N_COUNTERS, CLB_N_STATS, CLB_N_STAGES of include/columba_b200.h
COUNTER_NAMES = ["NODE_COUNTER", "TOTAL_REPORTED_POSITIONS", "CIGARS_IN_INDEX", "IN_TEXT_STARTED",
                 "ABORTED_IN_TEXT_VERIF", "CIGARS_IN_TEXT_VERIFICATION", "IMMEDIATE_SWITCH", "SEARCH_STARTED",
                 "DROPPED_UNIQUE_MATCHES", "NUMBER_OF_READS", "TOTAL_UNIQUE_MATCHES", "MAPPED_READS",
                 "TOTAL_UNIQUE_PAIRS", "MAPPED_PAIRS", "DISCORDANTLY_MAPPED_PAIRS", "MAPPED_HALF_PAIRS",
                 "UNPAIRED_BUT_MAPPED_PAIRS"]
STAT_NAMES = ["RANK_LOOKUPS", "LF_STEPS", "VERIFY_ROWS", "SEARCH_ROWS", "SEARCH_TASKS", "CANDIDATES", "FM_OCCS",
              "TEXT_OCCS", "SEARCH_LOOKUPS", "UNIQUE_CANDIDATES", "HEAVY_TASKS"]
STAGE_NAMES = ["prep", "partition", "search", "verify", "fm_expand", "text_filter"]
assert len(STAT_NAMES) == N_STATS and len(STAGE_NAMES) == N_STAGES
METRIC = {"hamming": 0, "edit": 1}
PARTITION = {"uniform": 0, "static": 1, "dynamic": 2}

OCC_DTYPE = np.dtype([("read", "<u4"), ("seq_id", "<u4"), ("pos", "<u4"), ("end", "<u4"), ("index_begin", "<u4"),
                      ("strand", "u1"), ("dist", "u1"), ("n_cigar", "u1"), ("trimmed", "u1")])


class ClbResult(C.Structure):
    _fields_ = [("n_reads", C.c_uint32), ("n_groups", C.c_uint32), ("n_occ", C.c_uint32), ("occ", C.c_void_p), ("cigar", C.c_void_p),
                ("cigar_stride", C.c_uint32), ("read_start", C.c_void_p), ("read_count", C.c_void_p),
                ("read_status", C.c_void_p), ("n_short_reads", C.c_uint32),
                ("counters", C.c_uint64 * N_COUNTERS), ("stats", C.c_uint64 * N_STATS),
                ("stage_ms", C.c_float * N_STAGES), ("total_ms", C.c_float), ("gpu_launches", C.c_uint32)]


class ClbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("columba_b200 error %d: %s" % (code, msg))
        self.code = code


EXPORTS = ["clb_abi_version", "clb_sizeof_result", "clb_device_count", "clb_index_load", "clb_index_free", "clb_index_text_length",
           "clb_index_num_seqs", "clb_index_seq_name", "clb_index_seq_start", "clb_index_device_bytes",
           "clb_ctx_create", "clb_ctx_free", "clb_last_error", "clb_ctx_set_strategy", "clb_ctx_set_options",
           "clb_ctx_set_stream", "clb_ctx_set_trim_max", "clb_ctx_set_filter_per_strand", "clb_ctx_supports_distance", "clb_ctx_num_parts", "clb_map_batch", "clb_map_batch_device", "clb_occ4",
           "clb_extend", "clb_find_sa", "clb_kmer_lookup", "clb_build_index", "clb_gather_ceiling", "clb_find_sa_static"]

STATIC_MAX_ROWS = 1 << 20  # CLB_STATIC_MAX_ROWS

_libs = {}


def load_library(path=None):
    path = os.path.abspath(path or DEFAULT_LIB)
    if path in _libs:
        return _libs[path]
    if not os.path.exists(path):
        raise ClbError(-3, "CUDA extension %s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)" % path)
    lib = C.CDLL(path)
    vp, u32p = C.c_void_p, C.POINTER(C.c_uint32)
    lib.clb_index_load.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
    lib.clb_index_free.argtypes = [vp]
    lib.clb_index_free.restype = None
    for f in ("clb_index_text_length", "clb_index_num_seqs"):
        getattr(lib, f).argtypes = [vp]
        getattr(lib, f).restype = C.c_uint32
    lib.clb_index_seq_name.argtypes = [vp, C.c_uint32]
    lib.clb_index_seq_name.restype = C.c_char_p
    lib.clb_index_seq_start.argtypes = [vp, C.c_uint32]
    lib.clb_index_seq_start.restype = C.c_uint32
    lib.clb_index_device_bytes.argtypes = [vp]
    lib.clb_index_device_bytes.restype = C.c_uint64
    lib.clb_ctx_create.argtypes = [vp, C.POINTER(vp)]
    lib.clb_ctx_free.argtypes = [vp]
    lib.clb_ctx_free.restype = None
    lib.clb_last_error.argtypes = [vp]
    lib.clb_last_error.restype = C.c_char_p
    lib.clb_ctx_set_strategy.argtypes = [vp, C.c_char_p, C.c_char_p, C.c_char_p]
    lib.clb_ctx_set_options.argtypes = [vp, C.c_int, C.c_int]
    lib.clb_ctx_set_stream.argtypes = [vp, vp]
    lib.clb_ctx_set_trim_max.argtypes = [vp, C.c_int]
    lib.clb_ctx_set_filter_per_strand.argtypes = [vp, C.c_int]
    lib.clb_ctx_supports_distance.argtypes = [vp, C.c_int]
    lib.clb_ctx_num_parts.argtypes = [vp, C.c_int]
    lib.clb_map_batch.argtypes = [vp, vp, vp, C.c_uint32, C.c_int, C.c_int, C.POINTER(ClbResult)]
    lib.clb_map_batch_device.argtypes = [vp, vp, vp, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int,
                                         C.POINTER(ClbResult)]
    lib.clb_occ4.argtypes = [vp, C.c_int, vp, C.c_uint32, vp]
    lib.clb_extend.argtypes = [vp, C.c_int, vp, C.c_uint32, vp]
    lib.clb_find_sa.argtypes = [vp, vp, C.c_uint32, vp]
    lib.clb_kmer_lookup.argtypes = [vp, vp, C.c_uint32, vp]
    lib.clb_build_index.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_size_t]
    if hasattr(lib, "clb_gather_ceiling"):  # CUDA library only (the CPU unit-test harness has no such kernel)
        lib.clb_gather_ceiling.argtypes = [vp, C.c_uint64, C.POINTER(C.c_float), C.POINTER(C.c_uint64)]
    if hasattr(lib, "clb_find_sa_static"):  # CUDA library only
        lib.clb_find_sa_static.argtypes = [vp, vp, C.c_uint32, vp]
    lib.clb_sizeof_result.restype = C.c_size_t
    if lib.clb_sizeof_result() != C.sizeof(ClbResult):
        raise ClbError(-6, "clb_result layout of %s differs from this binding's" % path)
    _libs[path] = lib
    return lib


def build_index(fasta, base, sa_sparse=4, lib_path=None):
    """Host-side index construction (reference file formats); not the hot path."""
    lib = load_library(lib_path)
    err = C.create_string_buffer(1024)
    rc = lib.clb_build_index(fasta.encode(), base.encode(), sa_sparse, err, 1024)
    if rc:
        raise ClbError(rc, err.value.decode())


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


CIG_OPS = "MID"


def cigar_string(ops):
    return "".join("%d%s" % (int(o) >> 2, CIG_OPS[int(o) & 3]) for o in ops)


class BatchResult:
    """Host copy of one clb_result."""

    def __init__(self, r, fetched=True):
        self.n_reads = r.n_reads
        self.n_groups = r.n_groups
        self.counters = dict(zip(COUNTER_NAMES, list(r.counters)))
        self.stats = dict(zip(STAT_NAMES, list(r.stats)))
        self.stage_ms = dict(zip(STAGE_NAMES, list(r.stage_ms)))
        self.total_ms = r.total_ms
        self.gpu_launches = r.gpu_launches
        self.cigar_stride = r.cigar_stride
        if fetched and r.n_reads:
            n = r.n_occ
            self.occ = np.frombuffer((C.c_char * (n * OCC_DTYPE.itemsize)).from_address(r.occ), dtype=OCC_DTYPE).copy() \
                if n else np.zeros(0, OCC_DTYPE)
            self.cigar = np.frombuffer((C.c_uint16 * (n * r.cigar_stride)).from_address(r.cigar), dtype=np.uint16) \
                .reshape(n, r.cigar_stride).copy() if n else np.zeros((0, r.cigar_stride), np.uint16)
            self.read_start = np.frombuffer((C.c_uint32 * r.n_groups).from_address(r.read_start), dtype=np.uint32).copy()
            self.read_count = np.frombuffer((C.c_uint32 * r.n_groups).from_address(r.read_count), dtype=np.uint32).copy()
            self.read_status = np.frombuffer((C.c_uint8 * r.n_reads).from_address(r.read_status), dtype=np.uint8).copy()
            self.n_short_reads = r.n_short_reads
        else:
            self.occ = None

    def tuples(self, read_names, seq_names):
        """Canonical (read, flag&16, seq, pos1, dist, CIGAR) tuples, sorted."""
        out = []
        for r in range(self.n_reads):
            s, c = int(self.read_start[r]), int(self.read_count[r])
            for i in range(s, s + c):
                o = self.occ[i]
                out.append((read_names[r], 16 if o["strand"] else 0, seq_names[int(o["seq_id"])], int(o["pos"]) + 1,
                            int(o["dist"]), cigar_string(self.cigar[i, :int(o["n_cigar"])]) if o["n_cigar"] else "*"))
        out.sort()
        return out


class Index:
    """Bidirectional FM-index resident in HBM (reference: FMIndex, src/fmindex/fmindex.h:403)."""

    def __init__(self, base, sa_sparse=4, kmer=10, in_text_switch=4, no_cigar=False, device=0, lib_path=None):
        self.lib = load_library(lib_path)
        h = C.c_void_p()
        rc = self.lib.clb_index_load(base.encode(), sa_sparse, kmer, in_text_switch, int(no_cigar), device,
                                     C.byref(h))
        if rc:
            raise ClbError(rc, self.lib.clb_last_error(None).decode())
        self.h = h
        self.text_length = self.lib.clb_index_text_length(h)
        n = self.lib.clb_index_num_seqs(h)
        self.seq_names = [self.lib.clb_index_seq_name(h, i).decode() for i in range(n)]
        self.seq_start = [self.lib.clb_index_seq_start(h, i) for i in range(n + 1)]
        self.device_bytes = self.lib.clb_index_device_bytes(h)

    def close(self):
        if self.h:
            self.lib.clb_index_free(self.h)
            self.h = None


class Mapper:
    """One mapping context (stream + scratch + strategy) on an Index.

    Mirrors the reference's SearchStrategy for ALL mode: `map_batch` is
    matchApprox (src/searchstrategy.h:2021) for a whole chunk of reads."""

    def __init__(self, index, scheme=("named", "columba"), metric="edit", partition="dynamic",
                 schemes_dir=SCHEMES_DIR):
        self.index = index
        self.lib = index.lib
        h = C.c_void_p()
        rc = self.lib.clb_ctx_create(index.h, C.byref(h))
        if rc:
            raise ClbError(rc, self.lib.clb_last_error(None).decode())
        self.h = h
        self._check(self.lib.clb_ctx_set_strategy(h, scheme[0].encode(), scheme[1].encode(), schemes_dir.encode()))
        self._check(self.lib.clb_ctx_set_options(h, METRIC[metric], PARTITION[partition]))

    def _check(self, rc):
        if rc:
            raise ClbError(rc, self.lib.clb_last_error(self.h).decode())

    def close(self):
        if self.h:
            self.lib.clb_ctx_free(self.h)
            self.h = None

    def set_stream(self, cuda_stream):
        """cuda_stream: integer cudaStream_t handle (0/None: the context's own stream)."""
        self._check(self.lib.clb_ctx_set_stream(self.h, C.c_void_p(cuda_stream or None)))

    def set_trim_max(self, largest_stratum):
        """largestStratum of findSeqName (0: the batch's max_dist); BEST mode sets its cut-off."""
        self._check(self.lib.clb_ctx_set_trim_max(self.h, int(largest_stratum)))

    def set_filter_per_strand(self, per_strand):
        """BEST mode filters each strand on its own (src/searchstrategy.cpp:777-811); results then come in
        2 * n_reads groups (2 * read + strand)."""
        self._check(self.lib.clb_ctx_set_filter_per_strand(self.h, int(bool(per_strand))))

    def supports(self, k):
        return bool(self.lib.clb_ctx_supports_distance(self.h, k))

    @staticmethod
    def pack(reads):
        """list of bytes/str -> (uint8 array, uint32 offsets)."""
        bs = [r.encode() if isinstance(r, str) else bytes(r) for r in reads]
        off = np.zeros(len(bs) + 1, np.uint32)
        off[1:] = np.cumsum([len(b) for b in bs])
        return np.frombuffer(b"".join(bs), dtype=np.uint8).copy(), off

    def map_batch(self, reads, offsets, max_dist, min_dist=0, allow_short=True):
        """reads: uint8 array of concatenated reads, offsets: uint32[n+1] (host)."""
        reads = np.ascontiguousarray(reads, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint32)
        r = ClbResult()
        rc = self.lib.clb_map_batch(self.h, _ptr(reads), _ptr(offsets), len(offsets) - 1, max_dist, min_dist,
                                    C.byref(r))
        self._check(rc)
        return BatchResult(r)

    def map_batch_device(self, d_reads_ptr, d_offsets_ptr, n_reads, max_read_len, max_dist, min_dist=0, fetch=False):
        r = ClbResult()
        self._check(self.lib.clb_map_batch_device(self.h, d_reads_ptr, d_offsets_ptr, n_reads, max_read_len,
                                                  max_dist, min_dist, int(fetch), C.byref(r)))
        return BatchResult(r, fetched=fetch)

    def gather_ceiling(self, n_lookups=1 << 28):
        """Measured random 32-byte gather rate on the loaded index: (lookups/s, GB/s)."""
        ms, done = C.c_float(), C.c_uint64()
        self._check(self.lib.clb_gather_ceiling(self.h, n_lookups, C.byref(ms), C.byref(done)))
        rate = done.value / (ms.value / 1e3)
        return rate, rate * 32 / 1e9

    # unit-level entry points
    def occ4(self, positions, reverse=False):
        p = np.ascontiguousarray(positions, np.uint32)
        out = np.zeros((len(p), 4), np.uint32)
        self._check(self.lib.clb_occ4(self.h, int(reverse), _ptr(p), len(p), _ptr(out)))
        return out

    def extend(self, parents, mode):
        p = np.ascontiguousarray(parents, np.uint32).reshape(-1, 4)
        out = np.zeros((len(p), 4, 4), np.uint32)
        self._check(self.lib.clb_extend(self.h, mode, _ptr(p), len(p), _ptr(out)))
        return out

    def find_sa(self, rows):
        p = np.ascontiguousarray(rows, np.uint32)
        out = np.zeros(len(p), np.uint32)
        self._check(self.lib.clb_find_sa(self.h, _ptr(p), len(p), _ptr(out)))
        return out

    def find_sa_static(self, rows):
        """clb_find_sa_static: find_sa staged in the library's static device array, no device allocation
        (at most STATIC_MAX_ROWS rows per call)."""
        p = np.ascontiguousarray(rows, np.uint32)
        out = np.zeros(len(p), np.uint32)
        self._check(self.lib.clb_find_sa_static(self.h, _ptr(p), len(p), _ptr(out)))
        return out

    def kmer_lookup(self, codes):
        p = np.ascontiguousarray(codes, np.uint32)
        out = np.zeros((len(p), 4), np.uint32)
        self._check(self.lib.clb_kmer_lookup(self.h, _ptr(p), len(p), _ptr(out)))
        return out
