import gzip
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
# several contexts on one GPU emulate a multi-GPU box in the sharded-table tests: give their streams hardware queues
# of their own (more streams than connections can serialise a spinning wait kernel ahead of the work it waits for)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
# the test queries are megabytes: share every table build out, however small (the library would count a small query on
# every GPU for itself); test_small_queries_are_not_shared_out turns the default back on
os.environ.setdefault("EREF_SHARD_MIN_BYTES", "0")
# ... and load every kernel when the library is loaded: a kernel's first launch may synchronise the context, which would
# wait for a wait kernel of ANOTHER emulated GPU spinning on this same device (on a real box each GPU has its own context)
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _have_gpu():
    try:
        from palace_b200 import eref
        return eref.load_library().eref_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    """ctypes view of oracle/liberef_oracle.so (the CPU restatement) — the checker, test-only."""
    from tests import oracle_util
    return oracle_util.load()


def load_golden(name):
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        meta = json.load(f)

    def rd(ext):
        with gzip.open(os.path.join(GOLDEN, f"{name}.{ext}.gz"), "rb") as g:
            return np.frombuffer(g.read(), np.uint8)

    return meta, rd("fa"), rd("r1.fq"), rd("r2.fq")


GOLDEN_CASES = ("basic", "short_reads", "edge")


@pytest.fixture(scope="session", params=GOLDEN_CASES)
def golden(request):
    return (request.param,) + load_golden(request.param)


def fq2_limit_inputs():
    """Inputs on which the reference's byte budget for fq2 bites (bin/extract_ref.cpp:1265-1291 hand file_size(fq1)
    to read_fastq for BOTH files, :942-945): derived from the committed "basic" fixture — fq1 keeps the first two thirds
    of its records, the rest is appended to fq2, so fq2 is twice as long as fq1 and the reference reads only its head.
    Expected outputs: tests/golden/fq2_limit.json (tests/golden/make_golden_fq2_limit.py, the unmodified reference)."""
    _, fa, f1, f2 = load_golden("basic")
    nl = np.flatnonzero(f1 == 10)
    cut = int(nl[4 * (nl.size // 4 * 2 // 3) - 1]) + 1      # a record boundary two thirds of the way through fq1
    with open(os.path.join(GOLDEN, "fq2_limit.json")) as f:
        meta = json.load(f)
    return meta, fa, f1[:cut].copy(), np.concatenate([f2, f1[cut:]])
