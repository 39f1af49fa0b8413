"""ctypes binding of libedsbwt_cuda.so (include/edsbwt.h) -- the only Python in the product path.

The directory name carries a hyphen (it mirrors the reference's name), so load it with
``importlib.import_module("eds-bwt_b200")``.  Nothing here computes: every call goes through the C ABI
into the CUDA library and raises if the library is missing or no GPU is present (no CPU fallback).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libedsbwt_cuda.so")

LOCATE_SA, LOCATE_WALK = 0, 1
LF_RANK, LF_MOVE = 0, 1
ENGINE_FUSED, ENGINE_LEVELSYNC, ENGINE_FUSED_STEP, ENGINE_ITEMS = 0, 1, 2, 3      # _FUSED: step-loop kernel for bytes, item kernel for 2-bit words
ROWS_GROUPED, ROWS_REFERENCE = 0, 1

OCC_DTYPE = np.dtype([("pat", "<u4"), ("str", "<u4"), ("seg", "<u4"), ("idx_in_seg", "<u4"), ("off", "<u4")])

EXPORTS = [
    "edsbwt_index_open", "edsbwt_index_open_ex", "edsbwt_index_info", "edsbwt_index_close", "edsbwt_search",
    "edsbwt_search_device", "edsbwt_result_sizes", "edsbwt_result_device_ptrs", "edsbwt_result_fetch",
    "edsbwt_result_free", "edsbwt_last_stats", "edsbwt_set_stats", "edsbwt_write_csv", "edsbwt_read_patterns",
    "edsbwt_build_from_eds", "edsbwt_build_from_collection", "edsbwt_gather_bench", "edsbwt_free",
    "edsbwt_strerror", "edsbwt_last_error", "edsbwt_version", "edsbwt_search_into", "edsbwt_host_alloc", "edsbwt_host_free",
    "edsbwt_set_chunk", "edsbwt_search_trace", "edsbwt_move_table_build", "edsbwt_move_table_fetch", "edsbwt_search_device_at", "edsbwt_gather_bench_ex",
    "edsbwt_set_row_order", "edsbwt_result_write_csv", "edsbwt_search_to_csv", "edsbwt_string_check", "edsbwt_search_into_fixed",
    "edsbwt_search_into_packed", "edsbwt_pack_patterns", "edsbwt_pack_patterns_alphabet",
    "edsbwt_multi_open", "edsbwt_multi_info", "edsbwt_multi_handle", "edsbwt_multi_close", "edsbwt_search_multi",
    "edsbwt_multi_result_info", "edsbwt_multi_result_device_ptrs", "edsbwt_multi_result_fetch", "edsbwt_multi_result_write_csv",
    "edsbwt_multi_result_free", "edsbwt_gather_bench_cfg", "edsbwt_move_table_save", "edsbwt_set_engine",
    "edsbwt_device_alloc", "edsbwt_device_free", "edsbwt_ipc_export", "edsbwt_ipc_open", "edsbwt_ipc_close", "edsbwt_result_push", "edsbwt_push_wait",
    "edsbwt_count_static",
]


class Info(C.Structure):
    _fields_ = [("n", C.c_uint32), ("n_text", C.c_uint32), ("n_seg", C.c_uint32), ("sigma", C.c_uint32),
                ("alphabet", C.c_uint8 * 8), ("table_occ", C.c_uint32 * 25), ("seed_k", C.c_uint32),
                ("locate_mode", C.c_uint32), ("device_bytes", C.c_uint64), ("seed_intervals", C.c_uint64),
                ("move_runs", C.c_uint32), ("move_intervals", C.c_uint32), ("move_a", C.c_uint32), ("move_rounds", C.c_uint32),
                ("from_file", C.c_uint32), ("engine", C.c_uint32), ("table_occ_wide", C.c_uint32 * 64)]


class GatherCfg(C.Structure):
    _fields_ = [("buffer_bytes", C.c_uint64), ("n_access", C.c_uint64), ("iters", C.c_int32), ("granule_bytes", C.c_int32), ("ilp", C.c_int32),
                ("ctas_per_sm", C.c_int32), ("dependent", C.c_int32), ("reserved", C.c_int32), ("mix_bytes", C.c_uint64), ("mix_frac", C.c_double)]


class OpenOpts(C.Structure):
    _fields_ = [("seed_k", C.c_int32), ("locate_mode", C.c_int32), ("verify", C.c_int32), ("build_move", C.c_int32),
                ("engine", C.c_int32), ("fused_cap", C.c_int32), ("fused_cap2", C.c_int32), ("move_a", C.c_int32), ("cache", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("I", C.c_uint64), ("S", C.c_uint64), ("E", C.c_uint64), ("L", C.c_uint64), ("O", C.c_uint64),
                ("steps", C.c_uint64), ("kernel_launches", C.c_uint64), ("lf_intervals", C.c_uint64),
                ("lf_kernel_ms", C.c_double), ("locate_kernel_ms", C.c_double), ("total_ms", C.c_double),
                ("frontier_peak", C.c_uint64), ("occ_blocks", C.c_uint64), ("overflow_patterns", C.c_uint64), ("levelsync_patterns", C.c_uint64),
                ("search_kernel_ms", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class EdsbwtError(RuntimeError):
    def __init__(self, code, detail):
        super().__init__("libedsbwt_cuda error %d: %s" % (code, detail))
        self.code = code


_lib = None


def lib():
    """Load the CUDA library; fail loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s is missing: run `python eds-bwt_b200/build.py` (nvcc, sm_100a). "
                              "There is no CPU fallback." % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.edsbwt_strerror.restype = C.c_char_p
        L.edsbwt_last_error.restype = C.c_char_p
        L.edsbwt_version.restype = C.c_char_p
        L.edsbwt_free.argtypes = [C.c_void_p]
        L.edsbwt_free.restype = None
        L.edsbwt_index_close.argtypes = [C.c_void_p]
        L.edsbwt_index_close.restype = None
        L.edsbwt_result_free.argtypes = [C.c_void_p]
        L.edsbwt_result_free.restype = None
        L.edsbwt_index_open_ex.argtypes = [C.c_char_p, C.c_int, C.POINTER(OpenOpts), C.POINTER(C.c_void_p)]
        L.edsbwt_index_info.argtypes = [C.c_void_p, C.POINTER(Info)]
        L.edsbwt_search.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_void_p,
                                    C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
        L.edsbwt_search_into.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_void_p,
                                         C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.edsbwt_search_into_fixed.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_void_p,
                                               C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.edsbwt_search_into_packed.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_void_p,
                                                C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.edsbwt_pack_patterns.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32)]
        L.edsbwt_pack_patterns_alphabet.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p,
                                                    C.POINTER(C.c_uint32)]
        L.edsbwt_multi_open.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.c_int, C.POINTER(OpenOpts), C.POINTER(C.c_void_p)]
        L.edsbwt_multi_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_int)]
        L.edsbwt_multi_handle.argtypes = [C.c_void_p, C.c_int]
        L.edsbwt_multi_handle.restype = C.c_void_p
        L.edsbwt_multi_close.argtypes = [C.c_void_p]
        L.edsbwt_multi_close.restype = None
        L.edsbwt_search_multi.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.edsbwt_multi_result_info.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.POINTER(C.c_double),
                                               C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
        L.edsbwt_multi_result_device_ptrs.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]
        L.edsbwt_multi_result_fetch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.edsbwt_multi_result_write_csv.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_uint64)]
        L.edsbwt_multi_result_free.argtypes = [C.c_void_p]
        L.edsbwt_multi_result_free.restype = None
        L.edsbwt_search_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]
        L.edsbwt_move_table_build.argtypes = [C.c_void_p]
        L.edsbwt_move_table_save.argtypes = [C.c_void_p, C.c_char_p]
        L.edsbwt_move_table_fetch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]
        L.edsbwt_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_uint64]
        L.edsbwt_host_free.argtypes = [C.c_void_p]
        L.edsbwt_host_free.restype = None
        L.edsbwt_set_chunk.argtypes = [C.c_void_p, C.c_uint32]
        L.edsbwt_search_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32,
                                           C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.edsbwt_search_device_at.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32,
                                              C.c_int, C.c_int, C.c_uint32, C.POINTER(C.c_void_p)]
        L.edsbwt_result_sizes.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)]
        L.edsbwt_result_device_ptrs.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.edsbwt_result_fetch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.edsbwt_last_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
        L.edsbwt_set_stats.argtypes = [C.c_void_p, C.c_int, C.c_uint32]
        L.edsbwt_write_csv.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_int, C.c_int]
        L.edsbwt_set_row_order.argtypes = [C.c_void_p, C.c_int]
        L.edsbwt_set_engine.argtypes = [C.c_void_p, C.c_int]
        L.edsbwt_device_alloc.argtypes = [C.c_int, C.c_uint64, C.POINTER(C.c_void_p)]
        L.edsbwt_device_free.argtypes = [C.c_int, C.c_void_p]
        L.edsbwt_device_free.restype = None
        L.edsbwt_ipc_export.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.edsbwt_ipc_open.argtypes = [C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]
        L.edsbwt_ipc_close.argtypes = [C.c_int, C.c_void_p]
        L.edsbwt_result_push.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.edsbwt_push_wait.argtypes = [C.c_void_p]
        L.edsbwt_result_write_csv.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_uint64)]
        L.edsbwt_search_to_csv.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_char_p, C.POINTER(C.c_uint64)]
        L.edsbwt_string_check.argtypes = [C.c_char_p, C.c_char_p]
        L.edsbwt_read_patterns.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_uint32)]
        L.edsbwt_build_from_eds.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.edsbwt_build_from_collection.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_char_p, C.c_int]
        L.edsbwt_gather_bench.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(C.c_double)]
        L.edsbwt_gather_bench_cfg.argtypes = [C.c_int, C.POINTER(GatherCfg), C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.edsbwt_gather_bench_ex.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.POINTER(C.c_double)]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise EdsbwtError(rc, lib().edsbwt_last_error().decode(errors="replace") or lib().edsbwt_strerror(rc).decode())


def pack_patterns(patterns):
    """list of bytes -> (uint8 array, uint64 offsets[npat+1])"""
    off = np.zeros(len(patterns) + 1, dtype=np.uint64)
    if len(patterns):
        off[1:] = np.cumsum([len(p) for p in patterns], dtype=np.uint64)
    data = np.frombuffer(b"".join(patterns), dtype=np.uint8).copy() if len(patterns) else np.zeros(0, np.uint8)
    if data.size == 0:
        data = np.zeros(1, np.uint8)
    return data, off


class Index:
    """Device-resident EDS-BWT index (edsbwt_index_open_ex)."""

    def __init__(self, base, device=0, seed_k=-1, locate_mode=LOCATE_SA, verify=True, build_move=False,
                 engine=ENGINE_FUSED, fused_cap=0, fused_cap2=0, move_a=0, cache=False):
        self._h = C.c_void_p()
        o = OpenOpts(seed_k=seed_k, locate_mode=locate_mode, verify=int(verify), build_move=int(build_move),
                     engine=engine, fused_cap=fused_cap, fused_cap2=fused_cap2, move_a=move_a, cache=int(cache))
        _check(lib().edsbwt_index_open_ex(os.fsencode(base), device, C.byref(o), C.byref(self._h)))
        self.device = device

    def close(self):
        if self._h:
            lib().edsbwt_index_close(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        i = Info()
        _check(lib().edsbwt_index_info(self._h, C.byref(i)))
        return i

    def set_stats(self, collect=True, canonical_k=12):
        _check(lib().edsbwt_set_stats(self._h, int(collect), canonical_k))

    def last_stats(self):
        s = Stats()
        _check(lib().edsbwt_last_stats(self._h, C.byref(s)))
        return s.as_dict()

    def search(self, patterns, locate=True, lf_variant=LF_RANK):
        """patterns: list of bytes, or (uint8 data, uint64 offsets).  -> (counts uint32[npat], occ structured array)"""
        data, off = pack_patterns(patterns) if isinstance(patterns, (list, tuple)) and not (
            len(patterns) == 2 and isinstance(patterns[0], np.ndarray)) else patterns
        data = np.ascontiguousarray(data, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.uint64)
        npat = len(off) - 1
        counts = np.zeros(max(npat, 1), dtype=np.uint32)
        occ_p = C.c_void_p()
        nocc = C.c_uint64(0)
        _check(lib().edsbwt_search(self._h, data.ctypes.data, off.ctypes.data, npat, int(locate), lf_variant,
                                   counts.ctypes.data, C.byref(occ_p), C.byref(nocc)))
        occ = np.zeros(0, dtype=OCC_DTYPE)
        if locate and nocc.value:
            buf = (C.c_uint8 * (nocc.value * OCC_DTYPE.itemsize)).from_address(occ_p.value)
            occ = np.frombuffer(buf, dtype=OCC_DTYPE).copy()
            lib().edsbwt_free(occ_p)
        return counts[:npat], occ

    def save_move_table(self, path=None):
        """edsbwt_move_table_save: <base>_MLF.edsbwt (or path); later opens of the same index files load it"""
        _check(lib().edsbwt_move_table_save(self._h, os.fsencode(path) if path else None))

    def move_table(self):
        """-> (p, q, dest, sym) columns of the move table (builds it when needed)."""
        r = C.c_uint32()
        _check(lib().edsbwt_move_table_fetch(self._h, None, None, None, None, 0, C.byref(r)))
        p = np.zeros(r.value, np.uint32); q = np.zeros(r.value, np.uint32); d = np.zeros(r.value, np.uint32); sy = np.zeros(r.value, np.uint8)
        _check(lib().edsbwt_move_table_fetch(self._h, p.ctypes.data, q.ctypes.data, d.ctypes.data, sy.ctypes.data, r.value, C.byref(r)))
        return p, q, d, sy

    def search_trace(self, patterns, lf_variant=LF_RANK):
        """count-only search + the reference's "interrupted in position" value per pattern (0 = no such line)."""
        data, off = pack_patterns(patterns)
        npat = len(off) - 1
        counts = np.zeros(max(npat, 1), dtype=np.uint32)
        pos = np.zeros(max(npat, 1), dtype=np.uint32)
        _check(lib().edsbwt_search_trace(self._h, data.ctypes.data, off.ctypes.data, npat, lf_variant, counts.ctypes.data, pos.ctypes.data))
        return counts[:npat], pos[:npat]

    def search_into(self, data, off, counts, occ=None, locate=True, lf_variant=LF_RANK):
        """Caller-owned host buffers (ideally from pinned_empty): counts uint32[npat], occ OCC_DTYPE[cap].
        -> number of occurrences; raises EdsbwtError(code -7) when occ is too small."""
        npat = len(off) - 1
        nocc = C.c_uint64(0)
        _check(lib().edsbwt_search_into(self._h, data.ctypes.data, off.ctypes.data, npat, int(locate), lf_variant,
                                        counts.ctypes.data, occ.ctypes.data if occ is not None else None,
                                        len(occ) if occ is not None else 0, C.byref(nocc)))
        return nocc.value

    def search_into_fixed(self, data, pat_len, counts, occ=None, locate=True, lf_variant=LF_RANK):
        """Equal-length patterns back to back in `data` (uint8, ideally pinned); no offset array crosses PCIe."""
        npat = len(data) // pat_len
        nocc = C.c_uint64(0)
        _check(lib().edsbwt_search_into_fixed(self._h, data.ctypes.data, pat_len, npat, int(locate), lf_variant, counts.ctypes.data,
                                              occ.ctypes.data if occ is not None else None, len(occ) if occ is not None else 0, C.byref(nocc)))
        return nocc.value

    def pack_patterns(self, data, pat_len, nthreads=1, words=None, invalid=None):
        """edsbwt_pack_patterns: equal-length pattern bytes -> (uint64 words, uint32 invalid bitmap, number flagged)."""
        npat = len(data) // pat_len
        stride = (pat_len + 31) // 32
        if words is None:
            words = np.zeros(max(1, npat * stride), np.uint64)
        if invalid is None:
            invalid = np.zeros(max(1, (npat + 31) // 32), np.uint32)
        nbad = C.c_uint32(0)
        _check(lib().edsbwt_pack_patterns(self._h, data.ctypes.data, pat_len, npat, nthreads, words.ctypes.data, invalid.ctypes.data, C.byref(nbad)))
        return words, invalid, nbad.value

    def search_into_packed(self, words, invalid, pat_len, npat, counts, occ=None, locate=True, lf_variant=LF_RANK):
        """The caller's own 2-bit words (layout: include/edsbwt.h); invalid may be None."""
        nocc = C.c_uint64(0)
        _check(lib().edsbwt_search_into_packed(self._h, words.ctypes.data, invalid.ctypes.data if invalid is not None else None, pat_len, npat,
                                               int(locate), lf_variant, counts.ctypes.data, occ.ctypes.data if occ is not None else None,
                                               len(occ) if occ is not None else 0, C.byref(nocc)))
        return nocc.value

    def push(self, result, d_counts_dst, d_occ_dst):
        """edsbwt_result_push: asynchronous copy of a DeviceResult's counts / tuples to device pointers (peer memory included)"""
        _check(lib().edsbwt_result_push(self._h, result._h, C.c_void_p(d_counts_dst) if d_counts_dst else None, C.c_void_p(d_occ_dst) if d_occ_dst else None))

    def push_wait(self):
        _check(lib().edsbwt_push_wait(self._h))

    def set_engine(self, engine):
        _check(lib().edsbwt_set_engine(self._h, engine))

    def set_row_order(self, reference=True):
        """ROWS_REFERENCE: tuples of a pattern in the reference's csv order."""
        _check(lib().edsbwt_set_row_order(self._h, ROWS_REFERENCE if reference else ROWS_GROUPED))

    def search_to_csv(self, patterns, csv_path, lf_variant=LF_RANK, header=True):
        """The reference constructor's whole loop: search + locate, rows in reference order formatted on the device.
        -> (counts, number of rows written)"""
        data, off = pack_patterns(patterns) if isinstance(patterns, (list, tuple)) and not (
            len(patterns) == 2 and isinstance(patterns[0], np.ndarray)) else patterns
        data = np.ascontiguousarray(data, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.uint64)
        npat = len(off) - 1
        counts = np.zeros(max(npat, 1), dtype=np.uint32)
        if header:
            write_csv(csv_path, np.zeros(0, dtype=OCC_DTYPE), move=(lf_variant == LF_MOVE))
        nocc = C.c_uint64(0)
        _check(lib().edsbwt_search_to_csv(self._h, data.ctypes.data, off.ctypes.data, npat, lf_variant, counts.ctypes.data,
                                          os.fsencode(csv_path), C.byref(nocc)))
        return counts[:npat], nocc.value

    def set_chunk(self, patterns_per_chunk):
        _check(lib().edsbwt_set_chunk(self._h, patterns_per_chunk))

    def search_device(self, d_bytes_ptr, d_off_ptr, npat, total_bytes, max_len, locate=True, lf_variant=LF_RANK, pat_base=0):
        """Device-pointer variant; returns a DeviceResult (results stay in HBM).  pat_base: index of the first pattern
        of this shard inside the whole batch (added to the pat field of the tuples)."""
        r = C.c_void_p()
        _check(lib().edsbwt_search_device_at(self._h, C.c_void_p(d_bytes_ptr), C.c_void_p(d_off_ptr), npat, total_bytes,
                                             max_len, int(locate), lf_variant, pat_base, C.byref(r)))
        return DeviceResult(r)

    def count_static(self, d_bytes_ptr, d_off_ptr, npat, d_counts_ptr):
        """edsbwt_count_static: count-only search that allocates nothing (scratch in a static device array).
        -> number of patterns left unresolved (0 on success); raises on any other error."""
        left = C.c_uint32()
        rc = lib().edsbwt_count_static(self._h, C.c_void_p(d_bytes_ptr), C.c_void_p(d_off_ptr), npat, C.c_void_p(d_counts_ptr), C.byref(left))
        if rc == -7 and left.value:          # EDSBWT_E_NOSPACE with a count: partial answer
            return left.value
        _check(rc)
        return 0


class MultiIndex:
    """The index replicated on several GPUs of this process (edsbwt_multi_open); patterns are sharded, the results gathered
    on the first device with NCCL while the next chunk is searched."""

    def __init__(self, base, devices=None, seed_k=-1, locate_mode=LOCATE_SA, verify=True, build_move=False, engine=ENGINE_FUSED,
                 fused_cap=0, fused_cap2=0, move_a=0):
        self._h = C.c_void_p()
        o = OpenOpts(seed_k=seed_k, locate_mode=locate_mode, verify=int(verify), build_move=int(build_move),
                     engine=engine, fused_cap=fused_cap, fused_cap2=fused_cap2, move_a=move_a, cache=0)
        devs = (C.c_int * len(devices))(*devices) if devices else None
        _check(lib().edsbwt_multi_open(os.fsencode(base), devs, len(devices) if devices else 0, C.byref(o), C.byref(self._h)))
        n, nccl = C.c_int(), C.c_int()
        d = (C.c_int * 64)()
        _check(lib().edsbwt_multi_info(self._h, C.byref(n), d, 64, C.byref(nccl)))
        self.devices, self.nccl = list(d[:n.value]), bool(nccl.value)

    def close(self):
        if self._h:
            lib().edsbwt_multi_close(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_row_order(self, reference=True):
        for i in range(len(self.devices)):
            _check(lib().edsbwt_set_row_order(lib().edsbwt_multi_handle(self._h, i), ROWS_REFERENCE if reference else ROWS_GROUPED))

    def search(self, patterns, locate=True, lf_variant=LF_RANK, chunks_per_gpu=0, fixed_len=0, csv_path=None, move=False):
        """patterns: list of bytes or (uint8 data, uint64 offsets); fixed_len > 0: data holds equal-length patterns, no offsets.
        -> (counts, occurrence tuples in ascending pattern order, info dict)"""
        if fixed_len:
            data = np.ascontiguousarray(patterns, dtype=np.uint8); off = None; npat = len(data) // fixed_len
        else:
            data, off = pack_patterns(patterns) if isinstance(patterns, (list, tuple)) and not (
                len(patterns) == 2 and isinstance(patterns[0], np.ndarray)) else patterns
            data = np.ascontiguousarray(data, dtype=np.uint8); off = np.ascontiguousarray(off, dtype=np.uint64); npat = len(off) - 1
        r = C.c_void_p()
        _check(lib().edsbwt_search_multi(self._h, data.ctypes.data, off.ctypes.data if off is not None else None, fixed_len, npat,
                                         int(locate), lf_variant, chunks_per_gpu, C.byref(r)))
        try:
            n, o, ns = C.c_uint32(), C.c_uint64(), C.c_uint32()
            sm, gm, wm, gb = C.c_double(), C.c_double(), C.c_double(), C.c_uint64()
            _check(lib().edsbwt_multi_result_info(r, C.byref(n), C.byref(o), C.byref(ns), C.byref(sm), C.byref(gm), C.byref(wm), C.byref(gb)))
            counts = np.zeros(max(n.value, 1), np.uint32); occ = np.zeros(max(o.value, 1), OCC_DTYPE)
            _check(lib().edsbwt_multi_result_fetch(r, counts.ctypes.data, occ.ctypes.data))
            if csv_path:
                _check(lib().edsbwt_multi_result_write_csv(self._h, r, os.fsencode(csv_path), int(move), 0, None))
            info = {"segments": ns.value, "search_ms": sm.value, "gather_ms": gm.value, "wall_ms": wm.value, "gather_bytes": gb.value}
        finally:
            lib().edsbwt_multi_result_free(r)
        return counts[:n.value], occ[:o.value], info


class DeviceResult:
    def __init__(self, h):
        self._h = h
        n, o = C.c_uint32(), C.c_uint64()
        _check(lib().edsbwt_result_sizes(h, C.byref(n), C.byref(o)))
        self.npat, self.nocc = n.value, o.value

    def device_ptrs(self):
        c, o = C.c_void_p(), C.c_void_p()
        _check(lib().edsbwt_result_device_ptrs(self._h, C.byref(c), C.byref(o)))
        return c.value, o.value

    def write_csv(self, index, path, move=False, append=False):
        """rows formatted to decimal text on the device; -> bytes of row text written"""
        nb = C.c_uint64(0)
        _check(lib().edsbwt_result_write_csv(index._h, self._h, os.fsencode(path), int(move), int(append), C.byref(nb)))
        return nb.value

    def fetch(self):
        counts = np.zeros(max(self.npat, 1), dtype=np.uint32)
        occ = np.zeros(max(self.nocc, 1), dtype=OCC_DTYPE)
        _check(lib().edsbwt_result_fetch(self._h, counts.ctypes.data, occ.ctypes.data))
        return counts[:self.npat], occ[:self.nocc]

    def free(self):
        if self._h:
            lib().edsbwt_result_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class _Pinned:
    def __init__(self, nbytes):
        self.p = C.c_void_p()
        _check(lib().edsbwt_host_alloc(C.byref(self.p), nbytes))

    def __del__(self):
        try:
            lib().edsbwt_host_free(self.p)
        except Exception:
            pass


def pinned_empty(n, dtype):
    """numpy array over page-locked host memory (edsbwt_host_alloc); freed with the array."""
    dtype = np.dtype(dtype)
    nbytes = max(1, int(n) * dtype.itemsize)
    owner = _Pinned(nbytes)
    buf = (C.c_uint8 * nbytes).from_address(owner.p.value)
    buf._owner = owner          # the ctypes buffer is the base of the array and keeps the allocation alive
    return np.frombuffer(buf, dtype=dtype, count=int(n))


def pack_patterns_alphabet(alphabet, data, pat_len, nthreads=1):
    """edsbwt_pack_patterns_alphabet (host-only, no GPU): alphabet = end-marker byte followed by the letters."""
    alphabet = np.frombuffer(bytes(alphabet), dtype=np.uint8)
    data = np.ascontiguousarray(data, dtype=np.uint8)
    npat = len(data) // pat_len
    words = np.zeros(max(1, npat * ((pat_len + 31) // 32)), np.uint64)
    invalid = np.zeros(max(1, (npat + 31) // 32), np.uint32)
    nbad = C.c_uint32(0)
    _check(lib().edsbwt_pack_patterns_alphabet(alphabet.ctypes.data, len(alphabet), data.ctypes.data, pat_len, npat, nthreads,
                                               words.ctypes.data, invalid.ctypes.data, C.byref(nbad)))
    return words, invalid, nbad.value


def device_alloc(device, nbytes):
    p = C.c_void_p()
    _check(lib().edsbwt_device_alloc(device, nbytes, C.byref(p)))
    return p.value


def device_free(device, ptr):
    lib().edsbwt_device_free(device, C.c_void_p(ptr))


def ipc_export(device, ptr):
    h = (C.c_uint8 * 64)()
    _check(lib().edsbwt_ipc_export(device, C.c_void_p(ptr), h))
    return bytes(h)


def ipc_open(device, handle):
    h = (C.c_uint8 * 64).from_buffer_copy(handle)
    p = C.c_void_p()
    _check(lib().edsbwt_ipc_open(device, h, C.byref(p)))
    return p.value


def ipc_close(device, ptr):
    _check(lib().edsbwt_ipc_close(device, C.c_void_p(ptr)))


def occ_rows(occ, pat_base=0):
    """structured occurrence array -> sorted list of csv tuples (pat+1, str, seg, idx_in_seg, off)."""
    return sorted(zip((occ["pat"] + pat_base + 1).tolist(), occ["str"].tolist(), occ["seg"].tolist(),
                      occ["idx_in_seg"].tolist(), occ["off"].tolist()))


def write_csv(path, occ, pat_base=0, move=False, append=False):
    occ = np.ascontiguousarray(occ, dtype=OCC_DTYPE)
    _check(lib().edsbwt_write_csv(os.fsencode(path), occ.ctypes.data if len(occ) else None, len(occ), pat_base, int(move), int(append)))


def read_patterns(path):
    b, o, n = C.c_void_p(), C.c_void_p(), C.c_uint32()
    _check(lib().edsbwt_read_patterns(os.fsencode(path), C.byref(b), C.byref(o), C.byref(n)))
    off = np.frombuffer((C.c_uint64 * (n.value + 1)).from_address(o.value), dtype=np.uint64).copy()
    total = int(off[-1])
    data = np.frombuffer((C.c_uint8 * max(total, 1)).from_address(b.value), dtype=np.uint8).copy()[:total]
    lib().edsbwt_free(b)
    lib().edsbwt_free(o)
    return data, off


def string_check(in_path, out_path):
    """stringCheck (stringCheck.cpp): bracket the non-degenerate stretches of an EDS text; host-only."""
    _check(lib().edsbwt_string_check(os.fsencode(in_path), os.fsencode(out_path)))


def build_from_eds(eds_path, out_base, device=0):
    _check(lib().edsbwt_build_from_eds(os.fsencode(eds_path), os.fsencode(out_base), device))


def build_from_collection(text, str_off, seg_first, out_base, device=0):
    text = np.ascontiguousarray(text, dtype=np.uint8)
    str_off = np.ascontiguousarray(str_off, dtype=np.uint64)
    seg_first = np.ascontiguousarray(seg_first, dtype=np.uint8)
    _check(lib().edsbwt_build_from_collection(text.ctypes.data if text.size else None, str_off.ctypes.data, len(seg_first),
                                              seg_first.ctypes.data, os.fsencode(out_base), device))


def gather_bench_cfg(device=0, buffer_bytes=4 << 30, n_access=1 << 27, iters=3, granule=64, ilp=4, ctas_per_sm=0, dependent=False,
                     mix_bytes=0, mix_frac=0.0):
    """-> (algorithmic GB/s, requests/s) of the random-gather microbenchmark in one configuration (csrc/gather_bench.cu)"""
    cfg = GatherCfg(buffer_bytes=buffer_bytes, n_access=n_access, iters=iters, granule_bytes=granule, ilp=ilp, ctas_per_sm=ctas_per_sm,
                    dependent=int(dependent), reserved=0, mix_bytes=mix_bytes, mix_frac=mix_frac)
    g, r = C.c_double(), C.c_double()
    _check(lib().edsbwt_gather_bench_cfg(device, C.byref(cfg), C.byref(g), C.byref(r)))
    return g.value, r.value


def gather_bench(device=0, buffer_bytes=4 << 30, n_access=1 << 26, iters=5, granule=64):
    g = C.c_double()
    _check(lib().edsbwt_gather_bench_ex(device, buffer_bytes, n_access, iters, granule, C.byref(g)))
    return g.value
