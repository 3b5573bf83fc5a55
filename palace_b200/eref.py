"""Host-side mirror of PALACE's `eref` over the C ABI of liberef_gpu.so (include/eref_gpu.h).

The reference has no library interface for this path — it is the process
`eref fq1 fq2 phagedb.fasta interval_out hit_ratio perfect_hit_ratio threads`
(bin/extract_ref.cpp:1221-1231, palace:475-477).  `run_eref()` takes the same seven arguments and
returns what that process writes to stdout; `ErefGPU` exposes the phases underneath it
(query table, reference framing, scan) for tests and for bench.py.

There is no CPU path here: if liberef_gpu.so is missing or no CUDA device is visible, every entry
point raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import (POINTER, Structure, byref, c_char_p, c_float, c_int, c_int16, c_size_t, c_uint,
                    c_uint32, c_uint64, c_void_p)

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# EREF_GPU_LIB: another build of the same library (the bounds-checked one, `make -C palace_b200/csrc bounds`)
LIB_PATH = os.environ.get("EREF_GPU_LIB") or os.path.join(_HERE, "liberef_gpu.so")
EREF_BIN = os.path.join(_HERE, "bin", "eref")

K = 32
FLAG_NO_PRUNE = 1
FLAG_DIRECT_COUNT = 2
FLAG_KEEP_RAND_STATE = 4
FLAG_LEVEL_POP = 8
FLAG_RESIDENT_TEXT = 16
TABLE_HANDLE_BYTES = 192
TABLE_HOOK = ctypes.CFUNCTYPE(c_int, c_void_p, c_void_p)
DOWN_SAMPLING_SIZE = 2_000_000_000                       # bin/extract_ref.cpp:1230

RESULT_DTYPE = np.dtype([("ordinal", "<u4"), ("frag", "<u4"), ("el", "<u4"), ("ref_len", "<u4")])

# every symbol include/eref_gpu.h declares
ABI_SYMBOLS = (
    "eref_last_error", "eref_device_count", "eref_create", "eref_destroy", "eref_set_stream",
    "eref_query_reset", "eref_query_add", "eref_query_add_device", "eref_query_seq_bytes",
    "eref_ref_set", "eref_ref_set_device", "eref_ref_count", "eref_scan", "eref_scan_device", "eref_scan_device_async",
    "eref_get_codes", "eref_ref_record", "eref_query_popcount", "eref_timings", "eref_format_line",
    "eref_run_host", "eref_ref_records", "eref_index_body", "eref_query_census", "eref_query_census_device",
    "eref_fastq_limit", "eref_sample_ratio", "eref_draw_keep", "eref_query_add_masked", "eref_query_add_masked_device",
    "eref_table_export", "eref_table_set_part", "eref_table_set_peer", "eref_table_set_hook", "eref_table_part_range",
    "eref_table_join", "eref_sync", "eref_run_results_device", "eref_run_host_keep", "eref_set_serial",
    "eref_query_reserve", "eref_lists_export", "eref_lists_set_peer", "eref_query_set_share",
    "eref_census_noalloc",
)
NOALLOC_MAX_BYTES = 256 << 20          # EREF_NOALLOC_MAX_BYTES


class PhaseTimes(Structure):
    _fields_ = [(n, c_float) for n in (
        "query_h2d_ms", "query_pack_ms", "query_bucket_ms", "query_count_ms", "ref_h2d_ms", "ref_frame_ms",
        "ref_pack_ms", "scan_summary_ms", "scan_sample_ms", "scan_cand_ms", "scan_probe_ms", "scan_window_ms",
        "scan_interval_ms", "scan_d2h_ms")] + [(n, c_uint64) for n in (
        "kernel_launches", "ref_bases", "ref_words", "ref_need_words", "query_seq_bytes", "query_codes")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class ErefError(RuntimeError):
    pass


_lib = None


def load_library() -> ctypes.CDLL:
    """dlopen liberef_gpu.so and bind the ABI.  Fails loudly when the extension was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ErefError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(make -C palace_b200/csrc). There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    lib.eref_last_error.restype = c_char_p
    lib.eref_device_count.restype = c_int
    lib.eref_create.argtypes = [POINTER(c_void_p), c_int, POINTER(c_int16), c_float, c_float, c_uint]
    lib.eref_destroy.argtypes = [c_void_p]
    lib.eref_destroy.restype = None
    lib.eref_set_stream.argtypes = [c_void_p, c_void_p]
    lib.eref_query_reset.argtypes = [c_void_p]
    lib.eref_query_add.argtypes = [c_void_p, c_void_p, c_size_t]
    lib.eref_query_add_device.argtypes = [c_void_p, c_void_p, c_size_t]
    lib.eref_query_seq_bytes.argtypes = [c_void_p, POINTER(c_uint64)]
    lib.eref_ref_set.argtypes = [c_void_p, c_void_p, c_size_t]
    lib.eref_ref_set_device.argtypes = [c_void_p, c_void_p, c_size_t]
    lib.eref_ref_count.argtypes = [c_void_p, POINTER(c_size_t), POINTER(c_size_t), POINTER(c_uint64)]
    lib.eref_scan.argtypes = [c_void_p, c_void_p, c_size_t, POINTER(c_size_t)]
    lib.eref_scan_device.argtypes = [c_void_p, POINTER(c_void_p), POINTER(c_size_t)]
    lib.eref_scan_device_async.argtypes = [c_void_p, POINTER(c_void_p), POINTER(c_size_t)]
    lib.eref_run_host.argtypes = [c_void_p, c_void_p, c_size_t, c_void_p, c_size_t, c_void_p, c_size_t, c_int, c_void_p,
                                  c_size_t, POINTER(c_size_t)]
    lib.eref_get_codes.argtypes = [c_void_p, c_uint32, c_void_p, c_size_t, POINTER(c_size_t)]
    lib.eref_ref_record.argtypes = [c_void_p, c_uint32, POINTER(c_uint64), POINTER(c_uint32), POINTER(c_uint32)]
    lib.eref_query_popcount.argtypes = [c_void_p, c_int, POINTER(c_uint64)]
    lib.eref_query_census.argtypes = [c_void_p, c_void_p, c_size_t, POINTER(c_uint64), POINTER(c_uint64)]
    lib.eref_query_census_device.argtypes = [c_void_p, c_void_p, c_size_t, POINTER(c_uint64), POINTER(c_uint64)]
    lib.eref_fastq_limit.argtypes = [c_void_p, c_size_t, c_uint64]
    lib.eref_fastq_limit.restype = c_size_t
    lib.eref_sample_ratio.argtypes = [c_uint64, c_uint64, POINTER(c_int)]
    lib.eref_draw_keep.argtypes = [c_int, c_uint64, c_void_p]
    lib.eref_draw_keep.restype = None
    lib.eref_query_add_masked.argtypes = [c_void_p, c_void_p, c_size_t, c_void_p, c_uint64]
    lib.eref_query_add_masked_device.argtypes = [c_void_p, c_void_p, c_size_t, c_void_p, c_uint64]
    lib.eref_ref_records.argtypes = [c_void_p, c_void_p, c_void_p, c_size_t, POINTER(c_size_t)]
    lib.eref_index_body.argtypes = [c_void_p, c_uint32, c_void_p, c_size_t, POINTER(c_size_t), POINTER(c_uint32)]
    lib.eref_table_export.argtypes = [c_void_p, c_void_p, POINTER(c_void_p), POINTER(c_void_p), POINTER(c_void_p)]
    lib.eref_table_set_part.argtypes = [c_void_p, c_int, c_int]
    lib.eref_table_set_peer.argtypes = [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]
    lib.eref_table_set_hook.argtypes = [c_void_p, TABLE_HOOK, c_void_p]
    lib.eref_table_part_range.argtypes = [c_void_p, c_int, POINTER(c_size_t), POINTER(c_size_t), POINTER(c_size_t), POINTER(c_size_t)]
    lib.eref_table_join.argtypes = [c_void_p]
    lib.eref_sync.argtypes = [c_void_p]
    lib.eref_set_serial.argtypes = [c_void_p, c_int]
    lib.eref_query_reserve.argtypes = [c_void_p, c_size_t, c_int]
    lib.eref_lists_export.argtypes = [c_void_p, c_void_p, POINTER(c_void_p)]
    lib.eref_lists_set_peer.argtypes = [c_void_p, c_int, c_void_p, c_void_p]
    lib.eref_query_set_share.argtypes = [c_void_p, c_int]
    lib.eref_run_results_device.argtypes = [c_void_p, POINTER(c_void_p), POINTER(c_size_t)]
    lib.eref_run_host_keep.argtypes = [c_void_p, c_void_p, c_uint64, c_void_p, c_uint64]
    lib.eref_timings.argtypes = [c_void_p, POINTER(PhaseTimes)]
    lib.eref_format_line.argtypes = [c_void_p, ctypes.c_char_p, c_size_t]
    lib.eref_census_noalloc.argtypes = [c_void_p, c_size_t, c_void_p, c_void_p]
    _lib = lib
    return lib


def _check(lib, rc, what):
    if rc != 0:
        raise ErefError(f"{what} failed ({rc}): {lib.eref_last_error().decode(errors='replace')}")


def census_noalloc(fastq_dev_ptr: int, n: int, out_dev_ptr: int, stream: int = 0):
    """eref_census_noalloc (no counterpart in the reference): the census of a FASTQ text in DEVICE memory on the current
    device, queued on `stream`, results (sequence-line bytes, sequence lines) written to two uint64 at out_dev_ptr.
    Allocates nothing; texts above NOALLOC_MAX_BYTES are refused."""
    lib = load_library()
    _check(lib, lib.eref_census_noalloc(c_void_p(fastq_dev_ptr), n, c_void_p(out_dev_ptr), c_void_p(stream)),
           "eref_census_noalloc")


# ---- coder assignment (bin/extract_ref.cpp:1082-1122) -------------------------------------------

_ORDERINGS = ((0, 1, 2), (0, 2, 1), (1, 2, 0), (1, 0, 2), (2, 0, 1), (2, 1, 0))   # permu[18], :1084


def random_coder(seed: int) -> np.ndarray:
    """random_coder (:1082-1102): glibc srand(seed), then 32 draws of rand() % 6."""
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(c_uint(seed))
    out = np.zeros(96, np.int16)
    for z in range(K):
        out[z * 3: z * 3 + 3] = _ORDERINGS[libc.rand() % 6]
    return out


def read_coder(index_path: str) -> np.ndarray:
    """saved_random_coder (:1104-1122): low 16 bits of the first 96 u32 words of <db>.k32.index.dat."""
    w = np.fromfile(index_path, dtype="<u4", count=96)
    if w.size != 96:
        raise ErefError(f"{index_path}: truncated index header")
    return (w & 0xFFFF).astype(np.int16)


def _as_bytes_array(buf) -> np.ndarray:
    if isinstance(buf, np.ndarray):
        a = buf if buf.dtype == np.uint8 else buf.view(np.uint8)
        return np.ascontiguousarray(a)
    return np.frombuffer(bytes(buf) if not isinstance(buf, (bytes, bytearray, memoryview)) else buf, dtype=np.uint8)


class DeviceBytes:
    """n bytes of text that already sit in device memory (eref_run_host takes either kind of pointer for the FASTQ files)"""

    def __init__(self, ptr: int, n: int):
        self.ptr, self.size = int(ptr), int(n)


class ErefGPU:
    """One context = one GPU: the query table, one reference text (whole DB or one shard), the scan."""

    def __init__(self, coder, hit_ratio: float = 0.9, perfect_hit_ratio: float = 0.85, device: int = 0,
                 prune: bool = True, bucketed: bool = True, level_pop: bool = False, resident_text: bool = False):
        self.lib = load_library()
        if self.lib.eref_device_count() <= 0:
            raise ErefError("no CUDA device visible; liberef_gpu.so has no CPU fallback")
        c = np.ascontiguousarray(np.asarray(coder, dtype=np.int16))
        if c.size != 96:
            raise ErefError("coder must have 96 entries")
        self._h = c_void_p()
        rc = self.lib.eref_create(byref(self._h), device, c.ctypes.data_as(POINTER(c_int16)),
                                  c_float(hit_ratio), c_float(perfect_hit_ratio),
                                  (0 if prune else FLAG_NO_PRUNE) | (0 if bucketed else FLAG_DIRECT_COUNT) |
                                  (FLAG_LEVEL_POP if level_pop else 0) | (FLAG_RESIDENT_TEXT if resident_text else 0))
        _check(self.lib, rc, "eref_create")
        self.device = device
        self._hook = None

    def close(self):
        if getattr(self, "_h", None):
            self.lib.eref_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_stream(self, cuda_stream: int):
        _check(self.lib, self.lib.eref_set_stream(self._h, c_void_p(cuda_stream)), "eref_set_stream")

    def query_reset(self):
        _check(self.lib, self.lib.eref_query_reset(self._h), "eref_query_reset")

    def query_add(self, fastq) -> int:
        """Count one FASTQ file held in HOST memory; returns its sequence-line byte sum (:1138-1140)."""
        a = _as_bytes_array(fastq)
        _check(self.lib, self.lib.eref_query_add(self._h, c_void_p(a.ctypes.data), a.size), "eref_query_add")
        return self.query_seq_bytes()

    def query_census(self, fastq):
        """cal_sam_ratio's pass (:1124-1148): (bytes on sequence lines, number of sequence lines)"""
        a = _as_bytes_array(fastq)
        b, l = c_uint64(), c_uint64()
        _check(self.lib, self.lib.eref_query_census(self._h, c_void_p(a.ctypes.data), a.size, byref(b), byref(l)),
               "eref_query_census")
        return b.value, l.value

    def query_add_masked(self, fastq, keep_bits: np.ndarray, n_lines: int) -> int:
        """read_fastq under down-sampling (:955-960): only the sequence lines whose bit is set are counted"""
        a = _as_bytes_array(fastq)
        k = np.ascontiguousarray(keep_bits, dtype="<u4")
        _check(self.lib, self.lib.eref_query_add_masked(self._h, c_void_p(a.ctypes.data), a.size, c_void_p(k.ctypes.data),
                                                        n_lines), "eref_query_add_masked")
        return self.query_seq_bytes()

    def query_add_device(self, ptr: int, n: int, want_bytes: bool = True):
        """want_bytes=False: do not wait for the file's kernels (the byte count of cal_sam_ratio is not asked for)"""
        _check(self.lib, self.lib.eref_query_add_device(self._h, c_void_p(ptr), n), "eref_query_add_device")
        return self.query_seq_bytes() if want_bytes else None

    def query_seq_bytes(self) -> int:
        v = c_uint64()
        _check(self.lib, self.lib.eref_query_seq_bytes(self._h, byref(v)), "eref_query_seq_bytes")
        return v.value

    def ref_set(self, fasta):
        a = _as_bytes_array(fasta)
        _check(self.lib, self.lib.eref_ref_set(self._h, c_void_p(a.ctypes.data), a.size), "eref_ref_set")

    def ref_set_device(self, ptr: int, n: int):
        _check(self.lib, self.lib.eref_ref_set_device(self._h, c_void_p(ptr), n), "eref_ref_set_device")

    def ref_count(self):
        k, r, b = c_size_t(), c_size_t(), c_uint64()
        _check(self.lib, self.lib.eref_ref_count(self._h, byref(k), byref(r), byref(b)), "eref_ref_count")
        return k.value, r.value, b.value

    def scan(self) -> np.ndarray:
        kept, _, _ = self.ref_count()
        out = np.zeros(max(kept, 1), RESULT_DTYPE)
        n = c_size_t()
        _check(self.lib, self.lib.eref_scan(self._h, c_void_p(out.ctypes.data), kept, byref(n)), "eref_scan")
        return out[:n.value]

    def run_host(self, fq1, fq2, fasta, chunks: int = 4, cap: int | None = None) -> np.ndarray:
        """The whole path from host buffers in one pipelined call (H2D overlapped with the kernels)."""
        a1 = fq1 if isinstance(fq1, DeviceBytes) else _as_bytes_array(fq1)
        a2 = fq2 if isinstance(fq2, DeviceBytes) else _as_bytes_array(fq2)
        p1 = a1.ptr if isinstance(a1, DeviceBytes) else a1.ctypes.data
        p2 = a2.ptr if isinstance(a2, DeviceBytes) else a2.ctypes.data
        if fasta is None:                                 # screen against the database already set (eref_ref_set)
            af = np.zeros(0, np.uint8)
            if cap is None:
                cap = self.ref_count()[0] + 1
        else:
            af = _as_bytes_array(fasta)
        if cap is None:
            cap = int(af.size // 33 + 16)                 # every kept record has more than 32 bases
        if getattr(self, "_run_out", None) is None or self._run_out.size < cap:
            self._run_out = np.zeros(cap, RESULT_DTYPE)
        n = c_size_t()
        rc = self.lib.eref_run_host(self._h, c_void_p(p1), a1.size, c_void_p(p2), a2.size,
                                    c_void_p(af.ctypes.data if fasta is not None else None), af.size, chunks,
                                    c_void_p(self._run_out.ctypes.data), cap,
                                    byref(n))
        _check(self.lib, rc, "eref_run_host")
        return self._run_out[:min(n.value, cap)]

    def run_results_device(self):
        """the records of the last run_host, still in HBM: (device pointer to n x 4 uint32, n)"""
        p, n = c_void_p(), c_size_t()
        _check(self.lib, self.lib.eref_run_results_device(self._h, byref(p), byref(n)), "eref_run_results_device")
        return p.value, n.value

    def run_host_keep(self, keep1: np.ndarray, lines1: int, keep2: np.ndarray, lines2: int):
        """down-sampling masks for the next run_host (drawn once for all GPUs that screen the same sample)"""
        k1, k2 = np.ascontiguousarray(keep1, dtype="<u4"), np.ascontiguousarray(keep2, dtype="<u4")
        _check(self.lib, self.lib.eref_run_host_keep(self._h, c_void_p(k1.ctypes.data), lines1, c_void_p(k2.ctypes.data),
                                                     lines2), "eref_run_host_keep")

    def scan_device(self, wait: bool = True):
        """Results stay in HBM: (device pointer to n x 4 uint32, n).  wait=False: valid in stream order on the
        context's stream (set_stream), nothing is waited for."""
        p, n = c_void_p(), c_size_t()
        fn = self.lib.eref_scan_device if wait else self.lib.eref_scan_device_async
        _check(self.lib, fn(self._h, byref(p), byref(n)), "eref_scan_device")
        return p.value, n.value

    def get_codes(self, ordinal: int) -> np.ndarray:
        n = c_size_t()
        _check(self.lib, self.lib.eref_get_codes(self._h, ordinal, None, 0, byref(n)), "eref_get_codes")
        out = np.zeros(n.value, "<u4")
        _check(self.lib, self.lib.eref_get_codes(self._h, ordinal, c_void_p(out.ctypes.data), out.size, byref(n)),
               "eref_get_codes")
        return out

    def ref_record(self, ordinal: int):
        off, hl, ln = c_uint64(), c_uint32(), c_uint32()
        _check(self.lib, self.lib.eref_ref_record(self._h, ordinal, byref(off), byref(hl), byref(ln)), "eref_ref_record")
        return off.value, hl.value, ln.value

    def ref_records(self):
        """(lengths, header offsets) of EVERY record of the reference text, kept or not, in file order"""
        n = c_size_t()
        _check(self.lib, self.lib.eref_ref_records(self._h, None, None, 0, byref(n)), "eref_ref_records")
        ln, off = np.zeros(n.value, "<u4"), np.zeros(n.value, "<u4")
        _check(self.lib, self.lib.eref_ref_records(self._h, c_void_p(ln.ctypes.data), c_void_p(off.ctypes.data), n.value,
                                                   byref(n)), "eref_ref_records")
        return ln, off

    def index_body(self, chunk_words: int = 1 << 26):
        """yield the body of <db>.k32.index.dat (:698, :711-738) in chunks of whole records"""
        kept = self.ref_count()[0]
        buf = np.zeros(chunk_words, "<u4")
        first = 1
        while first <= kept:
            n, nxt = c_size_t(), c_uint32()
            rc = self.lib.eref_index_body(self._h, first, c_void_p(buf.ctypes.data), buf.size, byref(n), byref(nxt))
            if rc == -4 and buf.size < (1 << 31):              # one record larger than the chunk: grow and retry
                buf = np.zeros(buf.size * 4, "<u4")
                continue
            _check(self.lib, rc, "eref_index_body")
            yield buf[:n.value].copy()
            first = nxt.value

    # ---- sharded table build (include/eref_gpu.h): GPU `part` of `parts` counts its slices of the table and stores them
    # into every peer's table over NVLink ----
    def table_export(self, ipc: bool = False):
        """(handles, bm3, summary, flags): IPC handle blob (bytes) for a context in another process, or the raw device
        pointers for one in this process"""
        h = ctypes.create_string_buffer(TABLE_HANDLE_BYTES) if ipc else None
        a, b, f = c_void_p(), c_void_p(), c_void_p()
        _check(self.lib, self.lib.eref_table_export(self._h, h, byref(a), byref(b), byref(f)), "eref_table_export")
        return (h.raw if ipc else None), a.value, b.value, f.value

    def table_set_part(self, part: int, parts: int):
        _check(self.lib, self.lib.eref_table_set_part(self._h, part, parts), "eref_table_set_part")

    def table_set_peer(self, peer_part: int, exported, peer_device: int = -1):
        h, a, b, f = exported
        _check(self.lib, self.lib.eref_table_set_peer(self._h, peer_part, peer_device, h, c_void_p(None if h else a),
                                                      c_void_p(None if h else b), c_void_p(None if h else f)),
               "eref_table_set_peer")

    # ---- shared reads: every GPU adds 1/N of each query file; a build pulls its table slices out of the peers' lists ----
    def query_reserve(self, max_file_bytes: int, max_files: int = 2):
        _check(self.lib, self.lib.eref_query_reserve(self._h, max_file_bytes, max_files), "eref_query_reserve")

    def lists_export(self, ipc: bool = False):
        h = ctypes.create_string_buffer(64) if ipc else None
        a = c_void_p()
        _check(self.lib, self.lib.eref_lists_export(self._h, h, byref(a)), "eref_lists_export")
        return (h.raw if ipc else None), a.value

    def lists_set_peer(self, peer_part: int, exported):
        h, a = exported
        _check(self.lib, self.lib.eref_lists_set_peer(self._h, peer_part, h, c_void_p(None if h else a)), "eref_lists_set_peer")

    def query_set_share(self, on: bool):
        _check(self.lib, self.lib.eref_query_set_share(self._h, 1 if on else 0), "eref_query_set_share")

    def table_set_hook(self, fn):
        """fn(cuda_stream: int) -> None completes the table on that stream (e.g. an in-place NCCL all-gather)"""
        def _cb(_arg, stream):
            try:
                fn(stream or 0)
                return 0
            except Exception:                              # an exception must not unwind through the C frame
                import traceback
                traceback.print_exc()
                return 1
        self._hook = TABLE_HOOK(_cb) if fn else TABLE_HOOK()
        _check(self.lib, self.lib.eref_table_set_hook(self._h, self._hook, None), "eref_table_set_hook")

    def table_part_range(self, part: int = -1):
        """(table offset, table bytes, summary offset, summary bytes) of a part (-1: this context's own)"""
        v = [c_size_t() for _ in range(4)]
        _check(self.lib, self.lib.eref_table_part_range(self._h, part, *[byref(x) for x in v]), "eref_table_part_range")
        return tuple(x.value for x in v)

    def table_join(self):
        _check(self.lib, self.lib.eref_table_join(self._h), "eref_table_join")

    def set_serial(self, on: bool):
        """profiling: no overlap between the context's streams, so phase times are per-kernel durations"""
        _check(self.lib, self.lib.eref_set_serial(self._h, 1 if on else 0), "eref_set_serial")

    def sync(self):
        _check(self.lib, self.lib.eref_sync(self._h), "eref_sync")

    def query_popcount(self, level: int) -> int:
        v = c_uint64()
        _check(self.lib, self.lib.eref_query_popcount(self._h, level, byref(v)), "eref_query_popcount")
        return v.value

    def timings(self) -> dict:
        t = PhaseTimes()
        _check(self.lib, self.lib.eref_timings(self._h, byref(t)), "eref_timings")
        return t.as_dict()


def record_name(header_line: bytes) -> str:
    """get_read_ID(line).substr(1) (:246-254, :689): the header up to the first '/', ' ' or tab, without '>'"""
    s = header_line
    for d in (b"/", b" ", b"\t"):
        k = s.find(d)
        if k >= 0:
            s = s[:k]
    return s[1:].decode("latin-1")


def write_index(fasta, coder, index_path: str, len_path: str | None = None, device: int = 0):
    """produce the reference's on-disk index for a FASTA text on the GPU —
    <db>.k32.index.dat (400-byte coder header :680-682, then per kept record its length and 3 codes per window
    start :698, :711-738) and <db>.genome.len.txt (name, record number, length, running length; :697, :800) —
    so that a CPU `eref` can share the DB directory.  Returns the number of u32 words written."""
    fasta = _as_bytes_array(fasta)
    c = np.asarray(coder, dtype=np.int64)
    hdr = np.zeros(100, "<u4")
    # the reference writes 4 bytes from a short array: word j = coder[j] | coder[j+1] << 16 (:680-682); only the low
    # halves are read back (:1104-1122)
    hdr[:96] = (c[:96] & 0xFFFF).astype("<u4")
    hdr[:95] |= ((c[1:96] & 0xFFFF) << 16).astype("<u4")
    words = 100
    with ErefGPU(coder, device=device) as g, open(index_path, "wb") as f:
        g.ref_set(fasta)
        f.write(hdr.tobytes())
        for chunk in g.index_body():
            f.write(chunk.tobytes())
            words += chunk.size
        if len_path:
            ln, off = g.ref_records()
            raw = fasta.tobytes()
            run = 0
            with open(len_path, "w") as lf:
                for r in range(ln.size):
                    run += int(ln[r])
                    if ln[r] > K:
                        if r == 0:
                            name = "start"
                        else:
                            e = raw.find(b"\n", int(off[r]))
                            name = record_name(raw[int(off[r]): e if e >= 0 else len(raw)])
                        lf.write(f"{name}\t{r}\t{int(ln[r])}\t{run}\n")
    return words


def format_results(results: np.ndarray, ordinal_base: int = 0) -> str:
    """The stdout of eref (:615-617) for an array of result records."""
    lib = load_library()
    buf = ctypes.create_string_buffer(256)
    out = []
    rec = np.zeros(1, RESULT_DTYPE)
    for r in results:
        rec[0] = r
        rec[0]["ordinal"] += ordinal_base
        k = lib.eref_format_line(c_void_p(rec.ctypes.data), buf, 256)
        if k > 0:
            out.append(buf.raw[:k].decode())
    return "".join(out)


def shard_offsets(fasta: np.ndarray, shards: int):
    """Byte offsets that split a FASTA text into `shards` ranges, each starting at a header line."""
    n = fasta.size
    cuts = [0]
    raw = fasta.tobytes() if shards > 1 else b""
    for s in range(1, shards):
        pos = max(cuts[-1], n // shards * s)
        q = raw.find(b"\n>", pos)
        cuts.append(n if q < 0 else q + 1)
    cuts.append(n)
    return cuts


def fastq_limit(fastq, end: int) -> int:
    """read_fastq's byte budget (:942-945): the length of the prefix of a FASTQ text the reference reads when main hands
    it the range [0, end).  main passes end = file_size(fq1) for BOTH files (:1265-1291): callers of the phase functions
    add fq2[:fastq_limit(fq2, fq1.size)] (run_eref and eref_run_host do it themselves).  Host arithmetic, no GPU."""
    a = _as_bytes_array(fastq)
    return int(load_library().eref_fastq_limit(c_void_p(a.ctypes.data), a.size, int(end)))


def sample_ratio(seq_bytes_fq1: int, down_sampling_size: int = DOWN_SAMPLING_SIZE) -> int:
    """cal_sam_ratio (:1124-1148): 100 * down_sampling_size / (2 * sequence bytes of fq1); below 100 the
    reference down-samples.  Raises where the reference divides by zero (empty fq1)."""
    lib = load_library()
    r = c_int()
    _check(lib, lib.eref_sample_ratio(seq_bytes_fq1, down_sampling_size, byref(r)), "eref_sample_ratio")
    return r.value


def draw_keep(ratio: int, n_lines: int, reseed: bool = True) -> np.ndarray:
    """The reference's per-read draws rand() % 100 < ratio (:955-957) as a bit mask, from libc's rand();
    reseed: srand(1) first, as main() does (:1236-1238)."""
    lib = load_library()
    if reseed:
        ctypes.CDLL("libc.so.6").srand(c_uint(1))
    bits = np.zeros((n_lines + 31) // 32 + 1, "<u4")
    lib.eref_draw_keep(ratio, n_lines, c_void_p(bits.ctypes.data))
    return bits


def run_eref(fq1, fq2, fasta, interval_out: str | None = None, hit_ratio: float = 0.9,
             perfect_hit_ratio: float = 0.85, threads: int = 1, *, coder=None, devices=(0,), prune: bool = True,
             bucketed: bool = True, pipelined: bool = False, chunks: int = 4, return_results: bool = False,
             down_sampling_size: int = DOWN_SAMPLING_SIZE, shard_table: bool = False, share_reads: bool = False):
    """The whole `eref` path on in-memory texts; returns the stdout text (and the raw records).

    `devices`: GPUs to shard the database over (one context each, run one after the other here —
    the concurrent multi-GPU drivers are palace_b200/bin/eref and bench.py under torchrun).
    """
    del threads                                            # a hint in the reference, unused here
    fq1, fq2 = _as_bytes_array(fq1), _as_bytes_array(fq2)
    fq2 = fq2[:fastq_limit(fq2, fq1.size)]                 # main reads fq2 only up to fq1's size (:1265-1291)
    if shard_table and len(devices) > 1:
        return _run_eref_sharded_table(fq1, fq2, fasta, interval_out, hit_ratio, perfect_hit_ratio, coder, devices, prune,
                                       pipelined, chunks, return_results, share_reads)
    if interval_out:
        open(interval_out, "w").close()                   # :825
    if coder is None:
        raise ErefError("run_eref needs the coder assignment (read_coder(index) or random_coder(seed))")
    fq1, fq2, fasta = _as_bytes_array(fq1), _as_bytes_array(fq2), _as_bytes_array(fasta)
    cuts = shard_offsets(fasta, len(devices))
    parts = []
    masks = None                                           # down-sampling draws: once, shared by every shard
    old_env = os.environ.get("EREF_DOWN_SAMPLING_SIZE")
    os.environ["EREF_DOWN_SAMPLING_SIZE"] = str(down_sampling_size)     # read by eref_create (eref_run_host's ratio)
    try:
        for i, dev in enumerate(devices):
            with ErefGPU(coder, hit_ratio, perfect_hit_ratio, dev, prune, bucketed) as g:
                if pipelined:
                    shard = fasta[cuts[i]:cuts[i + 1]]
                    parts.append(g.run_host(fq1, fq2, shard, chunks, cap=int(np.count_nonzero(shard == 62)) + 1).copy())
                    continue
                ratio = sample_ratio(g.query_add(fq1), down_sampling_size)
                if ratio < 100:                            # :955-960
                    if masks is None:
                        l1, l2 = g.query_census(fq1)[1], g.query_census(fq2)[1]
                        k1 = draw_keep(ratio, l1)
                        masks = (k1, l1, draw_keep(ratio, l2, reseed=False), l2)
                    g.query_reset()
                    g.query_add_masked(fq1, masks[0], masks[1])
                    g.query_add_masked(fq2, masks[2], masks[3])
                else:
                    g.query_add(fq2)
                g.ref_set(fasta[cuts[i]:cuts[i + 1]])
                parts.append(g.scan().copy())
    finally:
        if old_env is None:
            os.environ.pop("EREF_DOWN_SAMPLING_SIZE", None)
        else:
            os.environ["EREF_DOWN_SAMPLING_SIZE"] = old_env
    base = 0
    for p in parts:
        p["ordinal"] += base
        base += p.size
    res = np.concatenate(parts) if parts else np.zeros(0, RESULT_DTYPE)
    text = format_results(res)
    return (text, res) if return_results else text


def wire_table_peers(ctxs, share_file_bytes: int = 0, files: int = 2):
    """contexts of ONE process, one per part: every builder gets the raw table pointers of all the others; with
    share_file_bytes also the list arenas (shared reads: every context then adds 1/N of each query file)"""
    exported = [g.table_export() for g in ctxs]
    for i, g in enumerate(ctxs):
        g.table_set_part(i, len(ctxs))
    for i, g in enumerate(ctxs):
        for j, h in enumerate(ctxs):
            if i != j:
                g.table_set_peer(j, exported[j], h.device)
    if share_file_bytes:
        for g in ctxs:
            g.query_reserve(share_file_bytes, files)
        arenas = [g.lists_export() for g in ctxs]
        for i, g in enumerate(ctxs):
            for j in range(len(ctxs)):
                if i != j:
                    g.lists_set_peer(j, arenas[j])
            g.query_set_share(True)


def split_fastq(fq, shares: int):
    """a FASTQ text cut into `shares` pieces at 4-line record boundaries (each piece is a valid FASTQ text)"""
    a = _as_bytes_array(fq)
    nl = np.flatnonzero(a == 10)
    n_rec = (nl.size + (1 if a.size and a[-1] != 10 else 0)) // 4
    cuts = [0]
    for k in range(1, shares):
        r = n_rec * k // shares
        cuts.append(int(nl[4 * r - 1]) + 1 if r > 0 else 0)
    cuts.append(a.size)
    return [a[cuts[k]:cuts[k + 1]] for k in range(shares)]


def _run_eref_sharded_table(fq1, fq2, fasta, interval_out, hit_ratio, perfect_hit_ratio, coder, devices, prune, pipelined,
                            chunks, return_results, share_reads=False):
    """run_eref with the database AND the table build sharded over `devices` (entries may repeat: several contexts on
    one GPU emulate a multi-GPU box).  Context i counts its range of the 8192 table slices and stores them into
    every context's table; here the contexts run one after the other, so every build is waited for before any scan."""
    if interval_out:
        open(interval_out, "w").close()
    if coder is None:
        raise ErefError("run_eref needs the coder assignment (read_coder(index) or random_coder(seed))")
    fq1, fq2, fasta = _as_bytes_array(fq1), _as_bytes_array(fq2), _as_bytes_array(fasta)
    cuts = shard_offsets(fasta, len(devices))
    ctxs = [ErefGPU(coder, hit_ratio, perfect_hit_ratio, d, prune) for d in devices]
    try:
        if share_reads:
            # shared reads: context i adds the i-th of N pieces of each query file (cut at record boundaries)
            if pipelined:
                raise ErefError("eref_run_host takes whole query files: shared reads go through the phase calls")
            s1, s2 = split_fastq(fq1, len(ctxs)), split_fastq(fq2, len(ctxs))
            wire_table_peers(ctxs, share_file_bytes=max(max(x.size for x in s1), max(x.size for x in s2), 1))
            parts = [None] * len(ctxs)
            for i, g in enumerate(ctxs):
                g.query_add(s1[i]); g.query_add(s2[i])
            for i, g in enumerate(ctxs):
                g.ref_set(fasta[cuts[i]:cuts[i + 1]])      # queues this part's build: signal, wait, pull, count
            for g in ctxs:
                g.sync()
            for i, g in enumerate(ctxs):
                parts[i] = g.scan().copy()
            base = 0
            for p in parts:
                p["ordinal"] += base
                base += p.size
            res = np.concatenate(parts) if parts else np.zeros(0, RESULT_DTYPE)
            text = format_results(res)
            return (text, res) if return_results else text
        wire_table_peers(ctxs)
        parts = [None] * len(ctxs)
        if pipelined:
            # eref_run_host builds, joins and scans in one call, so the contexts must run side by side like the GPUs of
            # a box do: one host thread each (the calls release the GIL)
            import threading
            errs = []

            def work(i, g):
                try:
                    shard = fasta[cuts[i]:cuts[i + 1]]
                    parts[i] = g.run_host(fq1, fq2, shard, chunks, cap=int(np.count_nonzero(shard == 62)) + 1).copy()
                except Exception as e:                     # noqa: BLE001 - reported below
                    errs.append(e)
            ts = [threading.Thread(target=work, args=(i, g)) for i, g in enumerate(ctxs)]
            for t in ts:
                t.start()
            for t in ts:
                t.join()
            if errs:
                raise errs[0]
        else:
            for g in ctxs:
                if sample_ratio(g.query_add(fq1)) < 100:
                    raise ErefError("the sharded-table emulation does not down-sample")
                g.query_add(fq2)
            for i, g in enumerate(ctxs):
                g.ref_set(fasta[cuts[i]:cuts[i + 1]])      # queues this part's build beside the framing
            for g in ctxs:
                g.sync()                                   # every part has landed in every table
            for i, g in enumerate(ctxs):
                parts[i] = g.scan().copy()
    finally:
        for g in ctxs:
            g.close()
    base = 0
    for p in parts:
        p["ordinal"] += base
        base += p.size
    res = np.concatenate(parts) if parts else np.zeros(0, RESULT_DTYPE)
    text = format_results(res)
    return (text, res) if return_results else text
