"""ctypes binding of libpfbwt_b200.so (include/pfbwt_b200.h) used by tests/, bench.py and
__graft_entry__.py.  It is a thin view of the C ABI: no computation happens in Python and there
is no fallback - loading fails loudly when the library has not been built, and pfb_create fails
when no B200 is present."""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libpfbwt_b200.so")

EXPORTS = [
    "pfb_create", "pfb_destroy", "pfb_last_error", "pfb_create_error", "pfb_set_params", "pfb_add_fasta",
    "pfb_add_records", "pfb_finalize", "pfb_parse_bwt", "pfb_append", "pfb_save_parse", "pfb_load_parse",
    "pfb_build_bwt", "pfb_write_bwt", "pfb_get_sizes", "pfb_get_stats", "pfb_copy_dict", "pfb_copy_occ",
    "pfb_copy_parse", "pfb_copy_last", "pfb_copy_sai", "pfb_copy_bwlast", "pfb_copy_ilist", "pfb_copy_bwsai",
    "pfb_copy_bwt", "pfb_copy_sa", "pfb_copy_ssa", "pfb_copy_esa", "pfb_get_docs", "pfb_upload_text",
    "pfb_run_resident", "pfb_release_text", "pfb_clear",
    "pfb_nccl_unique_id", "pfb_shard_init_nccl", "pfb_shard_init_host", "pfb_get_shard_info", "pfb_get_pending_bases",
    "pfb_kat_trigger", "pfb_kat_trigger_static", "pfb_add_chunk", "pfb_add_fasta_range", "pfb_save_last_sai", "pfb_build_bwt_to_files", "pfb_build_workspace_bytes", "pfb_write_doc_array", "pfb_write_marker_array", "pfb_marker_join", "pfb_device_free_bytes", "pfb_save_parse_async", "pfb_wait_saves",
]


class ParseParams(C.Structure):
    _fields_ = [("w", C.c_uint32), ("p", C.c_uint64), ("get_sai", C.c_int), ("store_docs", C.c_int),
                ("verbose", C.c_int), ("trim_non_acgt", C.c_int), ("non_acgt_to_a", C.c_int)]


class Sizes(C.Structure):
    _fields_ = [("n", C.c_uint64), ("n_phrases", C.c_uint64), ("n_words", C.c_uint64), ("dict_bytes", C.c_uint64),
                ("r", C.c_uint64), ("easy", C.c_uint64), ("hard", C.c_uint64), ("bits", C.c_int),
                ("has_parse_bwt", C.c_int), ("has_bwt", C.c_int), ("has_sa", C.c_int), ("has_rssa", C.c_int)]


class Stats(C.Structure):
    _fields_ = [(k, C.c_float) for k in
                ["normalise_ms", "trigger_ms", "compact_ms", "fingerprint_ms", "dedup_ms", "phrase_sort_ms",
                 "dict_build_ms", "parse_sa_ms", "parse_bwt_ms", "dict_sort_ms", "groups_ms", "emit_ms", "runs_ms"]] + \
               [(k, C.c_uint32) for k in ["dedup_attempts", "phrase_sort_rounds", "parse_sa_rounds", "dict_sort_rounds"]] + \
               [("launches_own", C.c_uint64), ("launches_lib", C.c_uint64), ("dev_bytes_peak", C.c_uint64),
                ("radix_pass_ms", C.c_double), ("radix_passes", C.c_uint64), ("radix_pairs", C.c_uint64), ("run_ms", C.c_float),
                ("shard_exchange_ms", C.c_float), ("shard_merge_ms", C.c_float), ("shard_bytes_moved", C.c_uint64),
                ("ingest_ms", C.c_float), ("ingest_files", C.c_uint64),
                ("stream_buckets", C.c_uint64), ("stream_peak_bytes", C.c_uint64)]


class ShardInfo(C.Structure):
    _fields_ = [("rank", C.c_int), ("world", C.c_int)] + \
               [(k, C.c_uint64) for k in ["text_start", "bwt_off", "bwt_len", "ssa_off", "ssa_cnt", "esa_off", "esa_cnt"]]


# host all-gather-v callback of pfb_shard_init_host (include/pfbwt_b200.h)
ALLGATHERV = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64), C.c_int)


_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} not built - run `make` (or __graft_entry__.build()); there is no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    lib.pfb_create.restype = vp
    lib.pfb_create.argtypes = [C.c_int, C.c_int]
    lib.pfb_destroy.restype = None
    lib.pfb_destroy.argtypes = [vp]
    lib.pfb_last_error.restype = C.c_char_p
    lib.pfb_last_error.argtypes = [vp]
    lib.pfb_create_error.restype = C.c_char_p
    lib.pfb_set_params.argtypes = [vp, C.POINTER(ParseParams)]
    lib.pfb_add_fasta.argtypes = [vp, C.c_char_p]
    lib.pfb_add_records.argtypes = [vp, vp, vp, C.c_uint64]
    lib.pfb_add_chunk.argtypes = [vp, vp, C.c_uint64, C.c_int]
    lib.pfb_add_fasta_range.argtypes = [vp, C.c_char_p, C.c_uint64, C.c_uint64]
    for f in ["pfb_finalize", "pfb_parse_bwt", "pfb_upload_text", "pfb_release_text", "pfb_clear"]:
        getattr(lib, f).argtypes = [vp]
    lib.pfb_append.argtypes = [vp, vp]
    lib.pfb_kat_trigger.argtypes = [vp, vp, C.c_size_t, C.c_uint64, vp, vp]
    lib.pfb_kat_trigger_static.argtypes = [vp, vp, C.c_size_t, C.c_uint64, vp, vp]
    lib.pfb_save_parse.argtypes = [vp, C.c_char_p, C.c_int, C.c_int]
    lib.pfb_load_parse.argtypes = [vp, C.c_char_p, C.c_int, C.c_int]
    lib.pfb_save_last_sai.argtypes = [vp, C.c_char_p, C.c_int, C.c_int]
    lib.pfb_save_parse_async.argtypes = [vp, C.c_char_p, C.c_int, C.c_int]
    lib.pfb_wait_saves.argtypes = [vp]
    lib.pfb_build_bwt.argtypes = [vp, C.c_int, C.c_int]
    lib.pfb_run_resident.argtypes = [vp, C.c_int, C.c_int]
    lib.pfb_build_bwt_to_files.argtypes = [vp, C.c_char_p, C.c_int, C.c_int, C.c_uint64]
    lib.pfb_write_doc_array.argtypes = [vp, C.c_char_p]
    lib.pfb_write_marker_array.argtypes = [vp, C.c_char_p, C.c_char_p]
    lib.pfb_marker_join.argtypes = [vp, C.c_char_p, C.c_char_p, C.c_char_p]
    lib.pfb_device_free_bytes.argtypes = [vp, vp]
    lib.pfb_build_workspace_bytes.restype = C.c_uint64
    lib.pfb_build_workspace_bytes.argtypes = [vp, C.c_int, C.c_int]
    lib.pfb_write_bwt.argtypes = [vp, C.c_char_p, C.c_int, C.c_int, C.c_char_p]
    lib.pfb_get_sizes.argtypes = [vp, C.POINTER(Sizes)]
    lib.pfb_get_stats.argtypes = [vp, C.POINTER(Stats)]
    for f in ["dict", "occ", "parse", "last", "sai", "bwlast", "ilist", "bwsai", "bwt", "sa", "ssa", "esa"]:
        getattr(lib, "pfb_copy_" + f).argtypes = [vp, vp]
    lib.pfb_get_docs.restype = C.c_int64
    lib.pfb_get_docs.argtypes = [vp, vp]
    lib.pfb_nccl_unique_id.argtypes = [vp]
    lib.pfb_shard_init_nccl.argtypes = [vp, C.c_int, C.c_int, vp]
    lib.pfb_shard_init_host.argtypes = [vp, C.c_int, C.c_int, ALLGATHERV, vp]
    lib.pfb_get_shard_info.argtypes = [vp, C.POINTER(ShardInfo)]
    _lib = lib
    return lib


def nccl_unique_id():
    """128-byte NCCL id (call on rank 0, hand to the other ranks)"""
    lib = load()
    buf = C.create_string_buffer(128)
    if lib.pfb_nccl_unique_id(C.cast(buf, C.c_void_p)) != 0:
        raise RuntimeError("ncclGetUniqueId failed: " + lib.pfb_create_error().decode())
    return buf.raw


class PfbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"[{code}] {msg}")
        self.code = code
        self.msg = msg


class Context:
    """One pfb_ctx.  Method names follow the reference's classes (PfParser / PrefixFreeBWT)."""

    def __init__(self, device=0, bits=32, w=10, p=100, get_sai=False, non_acgt_to_a=False, store_docs=False):
        self.lib = load()
        self.h = self.lib.pfb_create(device, bits)
        if not self.h:
            raise PfbError(-5, self.lib.pfb_create_error().decode())
        self.bits = bits
        self.utype = np.uint32 if bits == 32 else np.uint64
        pp = ParseParams(w, p, int(get_sai), int(store_docs), 0, 0, int(non_acgt_to_a))
        self._ck(self.lib.pfb_set_params(self.h, C.byref(pp)))

    def _ck(self, rc):
        if rc != 0:
            raise PfbError(rc, self.lib.pfb_last_error(self.h).decode(errors="replace"))

    def close(self):
        if self.h:
            self.lib.pfb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # one job over several GPUs: attach this context to a communicator (collective calls afterwards)
    def shard_init_nccl(self, rank, world, unique_id):
        """unique_id: 128 bytes from nccl_unique_id() on rank 0, broadcast by the launcher"""
        buf = C.create_string_buffer(bytes(unique_id), 128)
        self._ck(self.lib.pfb_shard_init_nccl(self.h, rank, world, C.cast(buf, C.c_void_p)))

    def shard_init_host(self, rank, world, callback):
        self._cb = callback       # keep the ctypes thunk alive
        self._ck(self.lib.pfb_shard_init_host(self.h, rank, world, callback, None))

    def kat_trigger(self, kmers, p):
        """(wang_hash(kmers), p | hash) computed by the device functions of the trigger kernel"""
        k = np.ascontiguousarray(kmers, dtype=np.uint64)
        h = np.empty(k.shape[0], dtype=np.uint64)
        d = np.empty(k.shape[0], dtype=np.uint8)
        self._ck(self.lib.pfb_kat_trigger(self.h, k.ctypes.data, k.shape[0], int(p), h.ctypes.data, d.ctypes.data))
        return h, d

    def kat_trigger_static(self, kmers, p):
        """kat_trigger without device allocation: scratch in a static device array, at most KAT_STATIC_MAX k-mers"""
        k = np.ascontiguousarray(kmers, dtype=np.uint64)
        h = np.empty(k.shape[0], dtype=np.uint64)
        d = np.empty(k.shape[0], dtype=np.uint8)
        self._ck(self.lib.pfb_kat_trigger_static(self.h, k.ctypes.data, k.shape[0], int(p), h.ctypes.data, d.ctypes.data))
        return h, d

    def shard_info(self):
        s = ShardInfo()
        self._ck(self.lib.pfb_get_shard_info(self.h, C.byref(s)))
        return s

    # parse step
    def add_fasta(self, path):
        self._ck(self.lib.pfb_add_fasta(self.h, path.encode()))

    def add_records(self, seq, rec_len):
        """seq: uint8 numpy array or an integer host address; rec_len: uint64 array"""
        rec_len = np.ascontiguousarray(rec_len, dtype=np.uint64)
        addr = seq if isinstance(seq, int) else seq.ctypes.data
        self._ck(self.lib.pfb_add_records(self.h, addr, rec_len.ctypes.data, len(rec_len)))

    def add_chunk(self, seq, length=None, ends_record=True):
        """a piece of one record (mid-record cut); seq: uint8 numpy array or an integer host address (+ length)"""
        addr = seq if isinstance(seq, int) else seq.ctypes.data
        n = int(length) if length is not None else int(seq.shape[0])
        self._ck(self.lib.pfb_add_chunk(self.h, addr, n, int(ends_record)))

    def add_fasta_range(self, path, begin, end):
        self._ck(self.lib.pfb_add_fasta_range(self.h, path.encode(), int(begin), int(end)))

    def finalize(self):
        self._ck(self.lib.pfb_finalize(self.h))

    def bwt_of_parse(self):
        self._ck(self.lib.pfb_parse_bwt(self.h))

    def append(self, rhs):
        self._ck(self.lib.pfb_append(self.h, rhs.h))

    def save_parser(self, prefix, with_parse_bwt=False, with_sai=False):
        self._ck(self.lib.pfb_save_parse(self.h, prefix.encode(), int(with_parse_bwt), int(with_sai)))

    def save_parser_async(self, prefix, with_parse_bwt=0, with_sai=False):
        self._ck(self.lib.pfb_save_parse_async(self.h, prefix.encode(), int(with_parse_bwt), int(with_sai)))

    def wait_saves(self):
        self._ck(self.lib.pfb_wait_saves(self.h))

    def save_last_sai(self, prefix, with_last=True, with_sai=True):
        self._ck(self.lib.pfb_save_last_sai(self.h, prefix.encode(), int(with_last), int(with_sai)))

    def load_parser(self, prefix, with_parse_bwt=False, with_sai=False):
        self._ck(self.lib.pfb_load_parse(self.h, prefix.encode(), int(with_parse_bwt), int(with_sai)))

    # BWT build
    def generate_bwt(self, sa=False, rssa=False):
        self._ck(self.lib.pfb_build_bwt(self.h, int(sa), int(rssa)))

    def generate_bwt_to_files(self, prefix, sa=False, rssa=False, hbm_limit=0):
        """-m mode: build key range by key range, streaming every slice to <prefix>.bwt/.sa/.ssa/.esa"""
        self._ck(self.lib.pfb_build_bwt_to_files(self.h, prefix.encode(), int(sa), int(rssa), int(hbm_limit)))

    def build_workspace_bytes(self, sa=False, rssa=False):
        return int(self.lib.pfb_build_workspace_bytes(self.h, int(sa), int(rssa)))

    def write_doc_array(self, prefix):
        self._ck(self.lib.pfb_write_doc_array(self.h, prefix.encode()))

    def write_marker_array(self, mps_path, out_path):
        """mps_to_ma with the SA that generate_bwt left on the device"""
        self._ck(self.lib.pfb_write_marker_array(self.h, mps_path.encode(), out_path.encode()))

    def marker_join(self, mps_path, sa_path, out_path):
        """mps_to_ma: SA (u64 entries) from a file or '-' (stdin)"""
        self._ck(self.lib.pfb_marker_join(self.h, mps_path.encode(), sa_path.encode(), out_path.encode()))

    def write_bwt(self, prefix, sa=False, rssa=False):
        self._ck(self.lib.pfb_write_bwt(self.h, prefix.encode(), int(sa), int(rssa), None))

    # measurement hooks
    def upload_text(self):
        self._ck(self.lib.pfb_upload_text(self.h))

    def run_resident(self, sa=False, rssa=False):
        self._ck(self.lib.pfb_run_resident(self.h, int(sa), int(rssa)))

    def release_text(self):
        self._ck(self.lib.pfb_release_text(self.h))

    def clear(self):
        self._ck(self.lib.pfb_clear(self.h))

    # getters
    def sizes(self):
        s = Sizes()
        self._ck(self.lib.pfb_get_sizes(self.h, C.byref(s)))
        return s

    def stats(self):
        s = Stats()
        self._ck(self.lib.pfb_get_stats(self.h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in Stats._fields_}

    def _copy(self, what, count, dtype, out=None):
        a = out if out is not None else np.empty(count, dtype=dtype)
        self._ck(getattr(self.lib, "pfb_copy_" + what)(self.h, a.ctypes.data))
        return a

    def dict(self): return self._copy("dict", self.sizes().dict_bytes, np.uint8)
    def occ(self): return self._copy("occ", self.sizes().n_words, self.utype)
    def parse(self): return self._copy("parse", self.sizes().n_phrases, np.uint32)
    def last(self): return self._copy("last", self.sizes().n_phrases, np.uint8)
    def sai(self): return self._copy("sai", self.sizes().n_phrases, self.utype)
    def bwlast(self): return self._copy("bwlast", self.sizes().n_phrases + 1, np.uint8)
    def ilist(self): return self._copy("ilist", self.sizes().n_phrases + 1, self.utype)
    def bwsai(self): return self._copy("bwsai", self.sizes().n_phrases + 1, self.utype)
    # a rank of a multi-GPU job gets its slice (shard_info()); otherwise the slice is everything
    def bwt(self, out=None): return self._copy("bwt", self.shard_info().bwt_len, np.uint8, out)
    def sa(self, out=None): return self._copy("sa", self.shard_info().bwt_len, self.utype, out)
    def ssa(self): return self._copy("ssa", 2 * self.shard_info().ssa_cnt, self.utype)
    def esa(self): return self._copy("esa", 2 * self.shard_info().esa_cnt, self.utype)

    def docs(self):
        n = self.lib.pfb_get_docs(self.h, None)
        a = np.empty(max(n, 0), dtype=np.uint64)
        if n > 0:
            self.lib.pfb_get_docs(self.h, a.ctypes.data)
        return a
