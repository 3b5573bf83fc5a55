"""eref_census_noalloc (include/eref_gpu.h; no counterpart in the reference):
the FASTQ census with its scratch in a fixed __device__ array, against the existing entry points
eref_query_census_device / eref_query_census on the same texts.  Integer results: the tolerance is zero, as everywhere
in this suite.  Collected last on purpose; passes alone in a fresh process too."""
from ctypes import byref, c_uint64, c_void_p

import numpy as np
import pytest

from palace_b200 import eref, synth

pytestmark = pytest.mark.gpu


def _texts():
    db = synth.make_db(40, 1_000_000, 4242)
    big, other, _ = synth.make_reads(db, 1_500_000, 4243)
    big, other = np.array(big), np.array(other)
    small = other[: int(np.flatnonzero(other == 10)[4 * 57 - 1]) + 1]        # 57 whole records: far fewer tiles than `big`
    ragged = np.frombuffer(b"@a\nACGTNNAC\n+\n!!!!!!!!\n@b\n\n+\n\n@c\nacgtacgtacgtacgtacgtacgtacgtacgtacgt", np.uint8)  # ends inside a sequence line
    return {"big": big, "small": small, "ragged": ragged, "empty": np.zeros(0, np.uint8)}


def _python_census(a: np.ndarray):
    """cal_sam_ratio restated in three lines (bin/extract_ref.cpp:1134-1146): getline's lines, index % 4 == 1"""
    lines = a.tobytes().split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()                                 # getline yields nothing after a final newline
    seq = lines[1::4]
    return sum(len(x) for x in seq), len(seq)


def _existing_device(g, ptr, n):
    b, l = c_uint64(), c_uint64()
    eref._check(g.lib, g.lib.eref_query_census_device(g._h, c_void_p(ptr), n, byref(b), byref(l)), "eref_query_census_device")
    return b.value, l.value


def test_census_noalloc_matches_existing_entry_point():
    import torch
    torch.cuda.set_device(0)
    texts = _texts()
    dev = {}
    for k, a in texts.items():                      # 16 spare bytes: an empty tensor would have no address
        t = torch.zeros(a.size + 16, dtype=torch.uint8, device="cuda")
        t[: a.size] = torch.from_numpy(a.copy())
        dev[k] = t
    out = torch.empty(2, dtype=torch.int64, device="cuda")
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()

    def noalloc(k):
        out.fill_(-7)                                # the entry point must clear what has to start at zero
        torch.cuda.synchronize()
        eref.census_noalloc(dev[k].data_ptr(), texts[k].size, out.data_ptr(), stream.cuda_stream)
        stream.synchronize()
        b, l = out.cpu().tolist()
        return b, l

    with eref.ErefGPU(eref.random_coder(5)) as g:
        want = {}
        for k, a in texts.items():
            want[k] = _existing_device(g, dev[k].data_ptr(), a.size)
            assert want[k] == g.query_census(a) == _python_census(a), k
        assert want["big"][1] > 1000 and want["small"][1] == 57 and want["ragged"] == (8 + 0 + 36, 3)

        # two calls in a row, the second on a smaller text: the array still holds the first call's tiles
        assert noalloc("big") == want["big"]
        assert noalloc("small") == want["small"]
        # ... the same text twice, and every edge case after a big one
        assert noalloc("small") == want["small"]
        for k in ("big", "ragged", "big", "empty", "big"):
            assert noalloc(k) == want[k], k

        # queued back to back on one stream without a host wait in between: ordered, so both are right
        out2 = torch.empty(2, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        eref.census_noalloc(dev["big"].data_ptr(), texts["big"].size, out.data_ptr(), stream.cuda_stream)
        eref.census_noalloc(dev["small"].data_ptr(), texts["small"].size, out2.data_ptr(), stream.cuda_stream)
        stream.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0, "device memory changed across two eref_census_noalloc calls"
        assert tuple(out.cpu().tolist()) == want["big"] and tuple(out2.cpu().tolist()) == want["small"]

        # the existing entry point still gives the same after all that (nothing of its state is shared)
        assert _existing_device(g, dev["big"].data_ptr(), texts["big"].size) == want["big"]


def test_census_noalloc_refuses_what_it_cannot_hold():
    import torch
    torch.cuda.set_device(0)
    buf = torch.zeros(64, dtype=torch.uint8, device="cuda")
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    with pytest.raises(eref.ErefError, match="EREF_NOALLOC_MAX_BYTES"):
        eref.census_noalloc(buf.data_ptr(), eref.NOALLOC_MAX_BYTES + 1, out.data_ptr())     # refused before any access
    with pytest.raises(eref.ErefError, match="null"):
        eref.census_noalloc(buf.data_ptr(), 8, 0)
    with pytest.raises(eref.ErefError, match="aligned"):
        eref.census_noalloc(buf.data_ptr() + 1, 8, out.data_ptr())
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [0, 0]
