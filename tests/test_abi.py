"""CPU suite, part 2: the drop-in boundary.  The C-ABI library loads and exports every symbol that
include/eref_gpu.h declares; without a GPU the product fails loudly instead of falling back."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

from palace_b200 import eref
from tests.conftest import ROOT, _have_gpu


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "eref_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(eref_[a-z_0-9]+)\s*\(", src)))


def test_header_and_binding_list_agree():
    assert _header_symbols() == sorted(eref.ABI_SYMBOLS)


def test_library_exports_every_declared_symbol():
    assert os.path.exists(eref.LIB_PATH), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(eref.LIB_PATH)
    for sym in _header_symbols():
        assert hasattr(lib, sym), f"liberef_gpu.so does not export {sym}"


def test_no_torch_or_cpp_types_in_the_abi():
    src = open(os.path.join(ROOT, "include", "eref_gpu.h")).read()
    assert "torch" not in src and "std::" not in src and "at::" not in src
    assert 'extern "C"' in src


def test_format_line_is_host_only_and_matches_reference_format():
    lib = eref.load_library()
    rec = np.zeros(1, eref.RESULT_DTYPE)
    rec[0] = (43, 1, 11930, 11931)
    assert eref.format_results(rec) == "ref_index\t43\t1\t11930\t11931\t0.999916\n"     # reference output, /tmp probe
    rec[0] = (7, 2, 75, 100)                      # ratio 0.75 is NOT > 0.75 (bin/extract_ref.cpp:616)
    assert eref.format_results(rec) == ""
    rec[0] = (7, 2, 76, 100)
    assert eref.format_results(rec) == "ref_index\t7\t2\t76\t100\t0.76\n"
    rec[0] = (9, 1, 146290, 146291)
    assert eref.format_results(rec) == "ref_index\t9\t1\t146290\t146291\t0.999993\n"
    rec[0] = (1, 1, 1000, 1000)
    assert eref.format_results(rec) == "ref_index\t1\t1\t1000\t1000\t1\n"


@pytest.mark.skipif(_have_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_without_a_gpu(tmp_path):
    with pytest.raises(eref.ErefError):
        eref.ErefGPU(eref.random_coder(1))
    (tmp_path / "a.fq").write_bytes(b"@r\nACGT\n+\nIIII\n")
    (tmp_path / "db.fa").write_bytes(b">x\n" + b"ACGT" * 20 + b"\n")
    if os.path.exists(eref.EREF_BIN):
        r = subprocess.run([eref.EREF_BIN, "a.fq", "a.fq", "db.fa", "t.txt", "0.9", "0.85", "1"], cwd=tmp_path,
                           capture_output=True)
        assert r.returncode != 0 and r.stdout == b""
        assert b"no CUDA device" in r.stderr


def test_eref_binary_usage_error():
    if not os.path.exists(eref.EREF_BIN):
        pytest.skip("eref binary not built")
    r = subprocess.run([eref.EREF_BIN, "only", "three", "args"], capture_output=True)
    assert r.returncode == 2 and r.stdout == b""


def test_product_does_not_reach_into_the_oracle():
    """the oracle is the checker, never the product: nothing under palace_b200/ or include/ names it"""
    bad = []
    for base in ("palace_b200", "include"):
        for dp, _, fns in os.walk(os.path.join(ROOT, base)):
            for fn in fns:
                if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                    txt = open(os.path.join(dp, fn), errors="replace").read()
                    if re.search(r"oracle/|eref_oracle|liberef_oracle|eo_[a-z]+\(", txt):
                        bad.append(os.path.join(dp, fn))
    assert not bad, bad


def test_library_exports_nothing_the_header_does_not_declare():
    """the other direction: every eref_* symbol liberef_gpu.so exports is declared (and documented) in include/eref_gpu.h"""
    assert os.path.exists(eref.LIB_PATH)
    r = subprocess.run(["nm", "-D", "--defined-only", eref.LIB_PATH], capture_output=True, text=True)
    if r.returncode != 0:
        pytest.skip("nm is not available")
    exported = {ln.split()[-1] for ln in r.stdout.splitlines() if ln.split() and ln.split()[-1].startswith("eref_")}
    assert exported == set(_header_symbols())
