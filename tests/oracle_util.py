"""ctypes binding of oracle/liberef_oracle.so — the CPU restatement used as the CHECKER in tests.
Test infrastructure only: nothing under palace_b200/ imports this."""
import ctypes
import os
import subprocess
from ctypes import POINTER, byref, c_char_p, c_float, c_int, c_int16, c_size_t, c_uint, c_uint32, c_uint64, c_void_p

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liberef_oracle.so")
RESULT_DTYPE = np.dtype([("ordinal", "<u4"), ("frag", "<u4"), ("el", "<u4"), ("ref_len", "<u4")])


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        lib.eo_random_coder.argtypes = [c_uint, POINTER(c_int16)]
        lib.eo_read_coder.argtypes = [c_char_p, POINTER(c_int16)]
        lib.eo_code3_naive.argtypes = [POINTER(c_int16), c_void_p, POINTER(c_uint32)]
        lib.eo_code3_fast.argtypes = [POINTER(c_int16), c_void_p, POINTER(c_uint32)]
        lib.eo_fastq_seq_bytes.argtypes = [c_void_p, c_size_t]
        lib.eo_fastq_seq_bytes.restype = ctypes.c_long
        lib.eo_parse_fasta.argtypes = [c_void_p, c_size_t, POINTER(c_void_p), POINTER(c_void_p), POINTER(c_size_t),
                                       POINTER(c_size_t)]
        lib.eo_record_codes.argtypes = [POINTER(c_int16), c_void_p, c_uint32, c_void_p]
        lib.eo_run.argtypes = [c_void_p, c_size_t, c_void_p, c_size_t, c_void_p, c_size_t, POINTER(c_int16), c_float,
                               c_float, c_int, POINTER(c_void_p), POINTER(c_size_t)]
        lib.eo_run_sampled.argtypes = [c_void_p, c_size_t, c_void_p, c_size_t, c_void_p, c_size_t, POINTER(c_int16), c_float,
                                       c_float, c_int, ctypes.c_long, POINTER(c_void_p), POINTER(c_size_t)]
        lib.eo_format_line.argtypes = [c_void_p, c_char_p, c_size_t]
        lib.eo_write_index.argtypes = [c_char_p, POINTER(c_int16), c_void_p, c_size_t]
        self.libc = ctypes.CDLL("libc.so.6")
        self.libc.free.argtypes = [c_void_p]

    @staticmethod
    def _coder(coder):
        c = np.ascontiguousarray(np.asarray(coder, np.int16))
        assert c.size == 96
        return c, c.ctypes.data_as(POINTER(c_int16))

    def random_coder(self, seed):
        out = np.zeros(96, np.int16)
        self.lib.eo_random_coder(seed, out.ctypes.data_as(POINTER(c_int16)))
        return out

    def code3(self, coder, window: bytes, naive=False):
        c, cp = self._coder(coder)
        out = np.zeros(3, np.uint32)
        buf = ctypes.create_string_buffer(bytes(window), 32)
        fn = self.lib.eo_code3_naive if naive else self.lib.eo_code3_fast
        ok = fn(cp, buf, out.ctypes.data_as(POINTER(c_uint32)))
        return ok, out

    def fastq_limit(self, text, end):
        """the prefix of a FASTQ text read_fastq processes for the byte range [0, end) (:942-945), literal getline loop"""
        a = np.ascontiguousarray(text)
        self.lib.eo_fastq_limit.restype = c_size_t
        self.lib.eo_fastq_limit.argtypes = [c_void_p, c_size_t, ctypes.c_long]
        return int(self.lib.eo_fastq_limit(c_void_p(a.ctypes.data), a.size, int(end)))

    def fastq_seq_bytes(self, text):
        a = np.ascontiguousarray(text)
        return self.lib.eo_fastq_seq_bytes(c_void_p(a.ctypes.data), a.size)

    def parse_fasta(self, text):
        """-> list of kept record sequences (bytes), number of records framed"""
        a = np.ascontiguousarray(text)
        seq, off, kept, recs = c_void_p(), c_void_p(), c_size_t(), c_size_t()
        rc = self.lib.eo_parse_fasta(c_void_p(a.ctypes.data), a.size, byref(seq), byref(off), byref(kept), byref(recs))
        assert rc == 0
        offs = np.ctypeslib.as_array(ctypes.cast(off, POINTER(c_uint64)), shape=(kept.value + 1,)).copy()
        total = int(offs[-1])
        raw = ctypes.string_at(seq, total) if total else b""
        self.libc.free(seq); self.libc.free(off)
        return [raw[int(offs[i]):int(offs[i + 1])] for i in range(kept.value)], recs.value

    def record_codes(self, coder, seq: bytes):
        c, cp = self._coder(coder)
        n = (len(seq) - 31) * 3
        out = np.zeros(n, np.uint32)
        self.lib.eo_record_codes(cp, seq, len(seq), c_void_p(out.ctypes.data))
        return out

    def run(self, fq1, fq2, fasta, coder, hit_ratio=0.9, perfect_hit_ratio=0.85, threads=1,
            down_sampling_size=2_000_000_000):
        c, cp = self._coder(coder)
        a1, a2, af = (np.ascontiguousarray(x) for x in (fq1, fq2, fasta))
        res, n = c_void_p(), c_size_t()
        rc = self.lib.eo_run_sampled(c_void_p(a1.ctypes.data), a1.size, c_void_p(a2.ctypes.data), a2.size,
                                     c_void_p(af.ctypes.data), af.size, cp, c_float(hit_ratio), c_float(perfect_hit_ratio),
                                     threads, down_sampling_size, byref(res), byref(n))
        assert rc == 0
        if n.value:
            arr = np.frombuffer(ctypes.string_at(res, n.value * 16), RESULT_DTYPE).copy()
        else:
            arr = np.zeros(0, RESULT_DTYPE)
        self.libc.free(res)
        return arr

    def format(self, results):
        buf = ctypes.create_string_buffer(256)
        out = []
        rec = np.zeros(1, RESULT_DTYPE)
        for r in results:
            rec[0] = r
            k = self.lib.eo_format_line(c_void_p(rec.ctypes.data), buf, 256)
            if k > 0:
                out.append(buf.raw[:k].decode())
        return "".join(out)

    def write_index(self, path, coder, fasta):
        c, cp = self._coder(coder)
        a = np.ascontiguousarray(fasta)
        return self.lib.eo_write_index(path.encode(), cp, c_void_p(a.ctypes.data), a.size)


_inst = None


def load() -> Oracle:
    global _inst
    if _inst is None:
        if not os.path.exists(LIB):
            subprocess.run(["make", "-C", ORACLE_DIR, "liberef_oracle.so"], check=True, capture_output=True)
        _inst = Oracle(ctypes.CDLL(LIB))
    return _inst
